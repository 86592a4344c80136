"""Pins the CPU oracle (oracle/) against outputs of the reference itself (tests/golden/, produced
by oracle/gen_golden.py importing /root/reference).  CPU only."""
import numpy as np
import pytest
import torch

from oracle import attacks_ref, defenses_ref, jpeg_ref, preprocess_ref, resnet50_ref, weights
from oracle.gen_golden_preprocess import SHAPES as PREPROCESS_SHAPES, make_image


def _tiny(golden_dir):
    g = np.load(golden_dir / "tiny_linear.npz")
    wt = torch.from_numpy(g["wt"])
    return g, (lambda z: z.flatten(1) @ wt)


def test_fgsm_and_pgd_match_reference_bitwise(golden_dir):
    g, model = _tiny(golden_dir)
    mean, std = weights.mean_std()
    x, y = torch.from_numpy(g["x"]), torch.from_numpy(g["y"])
    eps, alpha = float(g["eps"]), float(g["alpha"])
    assert np.array_equal(attacks_ref.fgsm(model, x, y, eps=eps, mean=mean, std=std).numpy(), g["fgsm"])
    out = attacks_ref.pgd_linf(model, x, y, eps=eps, alpha=alpha, steps=7, mean=mean, std=std)
    assert np.array_equal(out.numpy(), g["pgd_nostart"])
    noise = weights.pgd_noise(x.shape, eps, 5)
    out = attacks_ref.pgd_linf(model, x, y, eps=eps, alpha=alpha, steps=7, mean=mean, std=std, noise=noise)
    assert np.array_equal(out.numpy(), g["pgd_seed5"])


def test_cw_matches_reference(golden_dir):
    g, model = _tiny(golden_dir)
    mean, std = weights.mean_std()
    x, y, yt = torch.from_numpy(g["x"]), torch.from_numpy(g["y"]), torch.from_numpy(g["yt"])
    xa, s, _ = attacks_ref.cw_l2(model, x, y, mean=mean, std=std, c=2.0, kappa=0.1, steps=25, lr=0.05)
    assert np.array_equal(s.numpy(), g["cwu_s"])
    np.testing.assert_allclose(xa.numpy(), g["cwu_x"], rtol=0, atol=2e-6)   # explicit Adam vs torch.optim.Adam
    xa, s, _ = attacks_ref.cw_l2(model, x, y, mean=mean, std=std, c=2.0, kappa=0.0, steps=25, lr=0.05,
                                 targeted=True, y_target=yt)
    assert np.array_equal(s.numpy(), g["cwt_s"])
    np.testing.assert_allclose(xa.numpy(), g["cwt_x"], rtol=0, atol=2e-6)
    with pytest.raises(ValueError):
        attacks_ref.cw_l2(model, x, y, mean=mean, std=std, targeted=True)


def test_median_oracle_matches_scipy():
    """The median-smoothing defense has no counterpart in the reference (its smoothing is the 3x3 mean), so its oracle
    is pinned against an independent implementation: scipy's median filter with constant-zero borders."""
    from scipy import ndimage
    rng = np.random.default_rng(11)
    for shape in ((2, 3, 17, 23), (1, 1, 1, 1), (1, 2, 1, 9), (1, 1, 6, 2), (1, 3, 224, 224)):
        x = (rng.random(shape, dtype=np.float32) * 1.2 - 0.1).astype(np.float32)
        want = ndimage.median_filter(x, size=(1, 1, 3, 3), mode="constant", cval=0.0)
        assert np.array_equal(defenses_ref.median3x3(x), want)


def test_defenses_match_reference_bitwise(golden_dir):
    g = np.load(golden_dir / "defense_small.npz")
    assert np.array_equal(defenses_ref.smooth3x3(g["x"]), g["smooth"])
    assert np.array_equal(defenses_ref.quantize(g["x"], 16), g["quant16"])
    assert np.array_equal(defenses_ref.quantize(g["x"], 4), g["quant4"])
    assert np.array_equal(defenses_ref.defend_input(g["x"]), g["defend"])


JPEG_CASES = ["img224", "odd37x51", "rand17x16", "one_pixel", "q30_64x48", "q95_40x40", "saturated32", "q50_30x23"]


def jpeg_case(g, name):
    x = g["img224_u8"].astype(np.float32) / np.float32(255) if name == "img224" else g[name + "_x"]
    return x, int(g[name + "_q"]), g[name + "_out_u8"]


@pytest.mark.parametrize("name", JPEG_CASES)
def test_jpeg_roundtrip_matches_reference_bitwise(golden_dir, name):
    """The libjpeg pipeline restated in oracle/jpeg_ref.py against the reference's _jpeg_compress_batch on PIL
    (tests/golden/jpeg_cases.npz, oracle/gen_golden_jpeg.py): every decoded byte identical, all sizes / qualities."""
    x, q, want = jpeg_case(np.load(golden_dir / "jpeg_cases.npz"), name)
    got = jpeg_ref.jpeg_compress_batch(x, q)
    assert np.array_equal(got, want.astype(np.float32) / np.float32(255))


def test_defend_input_with_jpeg_matches_reference_bitwise(golden_dir):
    g = np.load(golden_dir / "jpeg_cases.npz")
    assert np.array_equal(defenses_ref.defend_input(g["defend_x"], use_jpeg=True, jpeg_quality=75), g["defend_jpeg_out"])


def test_jpeg_building_blocks():
    qy, qc = jpeg_ref.quant_tables(75)
    assert qy[0, 0] == 8 and qy[7, 7] == 50 and qc[0, 0] == 9 and qc[7, 7] == 50      # Annex K scaled by 50 %
    assert jpeg_ref.quant_tables(100)[0].max() == 1 and jpeg_ref.quant_tables(1)[0].max() == 255
    # a flat block survives the codec exactly (DC only), and the DCT pair is near-inverse without quantisation
    flat = np.full((1, 1, 8, 8), 77, dtype=np.int64)
    assert np.array_equal(jpeg_ref.idct_islow(jpeg_ref.fdct_islow(flat - 128) // 8), flat)
    rng = np.random.default_rng(0)
    blk = rng.integers(0, 256, size=(4, 4, 8, 8))
    rec = jpeg_ref.idct_islow(np.round(jpeg_ref.fdct_islow(blk - 128) / 8).astype(np.int64))
    assert np.abs(rec - blk).max() <= 2
    # fancy upsampling of a constant plane is the constant; grey RGB maps to Y = v, Cb = Cr = 128 and back
    assert np.array_equal(jpeg_ref.h2v2_fancy_upsample(np.full((3, 5), 9)), np.full((6, 10), 9))
    grey = np.repeat(np.arange(256, dtype=np.uint8)[:, None, None], 3, axis=2)
    y, cb, cr = jpeg_ref.rgb_to_ycc(grey)
    assert np.array_equal(y[:, 0], np.arange(256)) and (cb == 128).all() and (cr == 128).all()
    assert np.array_equal(jpeg_ref.ycc_to_rgb(y, cb, cr), grey)


@pytest.mark.parametrize("shape", PREPROCESS_SHAPES + ["example"])
def test_preprocess_matches_reference_bitwise(golden_dir, shape):
    """Resize(256) -> CenterCrop(224) -> ToTensor restated (PIL's fixed-point resampling) against the reference's own
    transform on PIL (tests/golden/preprocess_cases.npz): every byte equal, up- and down-scaling, odd sizes."""
    g = np.load(golden_dir / "preprocess_cases.npz")
    if shape == "example":
        img, want = g["example_crop_in"], g["example_crop_out"]
    else:
        img, want = make_image(*shape), g["out_%dx%d" % shape]
    got = preprocess_ref.load_image_tensor(img)
    assert got.shape == (1, 3, 224, 224)
    assert np.array_equal(got[0], want.astype(np.float32) / np.float32(255))


def test_preprocess_geometry_and_to_pil():
    assert preprocess_ref.resized_size(762, 1024) == (256, 344) and preprocess_ref.resized_size(500, 333) == (384, 256)
    assert preprocess_ref.crop_offsets(256, 341) == (16, 58)          # Python round(58.5) = 58 (half to even)
    assert preprocess_ref.crop_offsets(256, 343) == (16, 60)          # round(59.5) = 60
    b, k = preprocess_ref.precompute_coeffs(512, 256)
    assert k.shape[1] == 5 and int(k[7].sum()) == 1 << 22 and tuple(b[0]) == (0, 3)
    with pytest.raises(ValueError):
        preprocess_ref.resize_center_crop_u8(np.zeros((100, 1000, 3), np.uint8), 256, 300)
    x = np.array([0.0, 0.5, 0.999, 1.0, 254.5 / 255], dtype=np.float32).reshape(1, 1, 5).repeat(3, 0)
    assert preprocess_ref.to_pil_u8(x)[0, :, 0].tolist() == [0, 127, 254, 255, 254]


@pytest.fixture(scope="module")
def fmodel():
    return resnet50_ref.FunctionalResNet50(weights.seeded_resnet50(0).state_dict())


def test_functional_resnet_matches_reference(golden_dir, fmodel):
    g = np.load(golden_dir / "resnet_n2.npz")
    mean, std = weights.mean_std()
    x = weights.synthetic_images(4)[:2]
    y = torch.from_numpy(g["y"])
    grad, logits = attacks_ref.ce_input_grad(fmodel, x, y, mean, std)
    ref = torch.from_numpy(g["logits"])
    assert float((logits - ref).norm() / ref.norm()) < 1e-5
    assert torch.equal(logits.argmax(1), y)
    gref = torch.from_numpy(g["grad"])
    # the reference's own fp32 noise floor on this model is ~4e-3 (meta.json: ref_fp32_vs_fp64)
    assert float((grad - gref).norm() / gref.norm()) < 2e-2
    mask = gref.abs() > 1e-6 * gref.abs().max()
    assert float((grad.sign() == gref.sign())[mask].float().mean()) > 0.998


def test_detector_and_evaluate_defenses_match_reference(golden_meta, fmodel):
    mean, std = weights.mean_std()
    xs = weights.synthetic_images(4)
    det = golden_meta["detector"]
    feats = fmodel.layer3_features(attacks_ref.normalize(xs, mean, std))
    scores = defenses_ref.layer3_score(feats)
    np.testing.assert_allclose(scores.numpy(), np.array(det["scores"], dtype=np.float32), rtol=2e-5)
    thr = defenses_ref.calibrate_threshold(torch.tensor(det["scores"]), 0.95)
    assert abs(thr - det["threshold"]) < 1e-6
    x = xs[:1]
    y = fmodel(attacks_ref.normalize(x, mean, std)).argmax(1)
    out, _ = defenses_ref.evaluate_defenses(fmodel, x, y, mean, std, "fgsm", 4 / 255, 2 / 255, 10, det["threshold"])
    want = det["evaluate_defenses"]["fgsm_4"][0]
    assert {k: int(v) for k, v in out.items()} == want


# ---------------------------------------------------------------- the conditioned (non-identity BatchNorm) fixture model
@pytest.fixture(scope="module")
def cond_model():
    return weights.conditioned_resnet50(0)


def test_conditioned_model_has_nontrivial_batchnorm_everywhere(cond_model):
    """Every one of the 53 BatchNorms folds to a non-trivial scale AND shift (the seed-0 model folds to 1/sqrt(1+eps), 0)."""
    sd = cond_model.state_dict()
    n = 0
    for k in sd:
        if k.endswith("running_var"):
            p = k[:-len(".running_var")]
            scale = sd[p + ".weight"] / torch.sqrt(sd[k] + 1e-5)
            shift = sd[p + ".bias"] - sd[p + ".running_mean"] * scale
            assert float(scale.std()) > 1e-3 and float(shift.abs().mean()) > 1e-3, p
            n += 1
    assert n == 53
    # the committed statistics are what a fresh calibration pass computes (to fp32 noise)
    fresh = weights.conditioned_resnet50(0, stats_file="/nonexistent").state_dict()
    for k in sd:
        if "running_" in k:
            assert float((sd[k] - fresh[k]).abs().max() / fresh[k].abs().max()) < 1e-4, k


def test_functional_resnet_matches_reference_on_the_conditioned_model(golden_dir, cond_model):
    g = np.load(golden_dir / "cond_resnet.npz")
    meta = __import__("json").loads((golden_dir / "meta_cond.json").read_text())
    mean, std = weights.mean_std()
    x = weights.mixed_images(8, seed=77)[:4]
    y = torch.from_numpy(g["y"])
    fm = resnet50_ref.FunctionalResNet50(cond_model.state_dict())
    grad, logits = attacks_ref.ce_input_grad(fm, x, y, mean, std)
    ref = torch.from_numpy(g["logits"])
    assert float((logits - ref).norm() / ref.norm()) < 1e-5
    assert torch.equal(logits.argmax(1), y) and len(set(y.tolist())) > 1          # predictions differ between images
    gref = torch.from_numpy(g["grad"])
    assert float((grad - gref).norm() / gref.norm()) < 3 * meta["ref_fp32_vs_fp64"]["grad_rel"] + 1e-3


def test_bf16_emulator_without_rounding_is_the_reference_network(golden_dir, cond_model):
    """oracle/resnet50_bf16emu.py with emulate=False == torchvision in float64 (the reference's fp64 golden)."""
    from oracle import resnet50_bf16emu as emu
    g = np.load(golden_dir / "cond_resnet.npz")
    mean, std = weights.mean_std()
    x = weights.mixed_images(8, seed=77)[:2]
    y = torch.from_numpy(g["y"])[:2]
    e = emu.Bf16EmuResNet50(cond_model.state_dict(), emulate=False)
    grad, logits = e.input_grad(x, y, mean, std, inv_n=1 / 4)          # the golden gradient is the mean over 4 images
    l64, g64 = torch.from_numpy(g["logits64"])[:2], torch.from_numpy(g["grad64"])[:2]
    assert float((logits - l64).norm() / l64.norm()) < 2e-6            # BN folded in fp32, like the product
    assert float((grad - g64).norm() / g64.norm()) < 2e-3


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
def test_launch_by_launch_oracle_reproduces_the_autograd_emulator(cond_model, precision):
    """oracle/trace_ref.py (explicit per-launch forward / input-gradient formulas, used teacher-forced against the CUDA
    trace) chained free-running == the autograd-based emulator: same logits bit for bit, same gradient."""
    from oracle import resnet50_bf16emu as emu, trace_ref
    mean, std = weights.mean_std()
    x = weights.mixed_images(2, seed=5)
    e = emu.Bf16EmuResNet50(cond_model.state_dict(), emulate=precision == "bf16")
    y = e(x, mean, std).argmax(1)
    grad, logits = e.input_grad(x, y, mean, std)
    w = trace_ref.ManualResNet50(cond_model.state_dict(), precision).walk(x, mean, std, y, 1 / 2)
    assert float((w["logits"].flatten(1) - logits).abs().max()) < 1e-6
    assert float((w["grad"] - grad).norm() / grad.norm()) < 1e-6
    rel, bad = trace_ref.compare(w["grad"], grad, precision == "bf16")
    assert bad == 0
    assert trace_ref.canonical("bwd l2.0.c2/*") == "bwd l2.0.c2" and trace_ref.canonical("bwd l3.0.c1+ds/3") == "bwd l3.0.c1+ds"


def test_exact_constant_divisions_host_checker(tmp_path):
    """oracle/check_div9.c: the branch-free s/9 of the smoothing kernels against IEEE division for ALL 2^32 floats, and the
    k/255 of the JPEG second pass for all 256 k, with C's correctly rounded fmaf (the device-side twin is
    advb200_test_div9 in the GPU suite).  About half a minute on 8 cores."""
    import shutil
    import subprocess
    from pathlib import Path
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    exe = tmp_path / "check_div9"
    src = Path(__file__).resolve().parent.parent / "oracle" / "check_div9.c"
    subprocess.run(["gcc", "-O2", "-mfma", "-fopenmp", "-ffp-contract=off", str(src), "-o", str(exe), "-lm"], check=True)
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=1200)
    assert r.returncode == 0, r.stdout
    assert "mismatches 0 " in r.stdout and "k/255 mismatches 0" in r.stdout
