"""Parity tests proper (run with -m gpu on a B200): the CUDA path, called through the C ABI /
Python mirror, against the CPU oracle on the same seeded inputs and against the committed golden
vectors produced by the reference itself (tests/golden/, oracle/gen_golden.py).

Tolerances (north_star): integer/elementwise work bit-exact; logits 1e-4 (fp32 validation mode)
/ 1e-2 (bf16) relative; input gradients are compared against the reference's own fp32-vs-fp64
noise floor on this random-init model (tests/golden/meta.json) — the bounds
asserted here are the measured ones, not the aspirational 1e-4 / 1e-2 the model cannot support.
"""
import json

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import attacks_ref, defenses_ref, resnet50_ref, weights  # noqa: E402


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def advb200_mod():
    import advb200
    return advb200


@pytest.fixture(scope="module")
def torch_model(dev):
    # a torch-on-GPU checker must not silently run TF32 convolutions (dtype control)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    return weights.seeded_resnet50(0).to(dev)


@pytest.fixture(scope="module")
def native(advb200_mod, torch_model):
    return {p: advb200_mod.NativeResNet50(torch_model, precision=p) for p in ("fp32", "bf16")}


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / b.norm())


# ------------------------------------------------------------------ update / defense kernels
def test_native_library_is_the_one_loaded(advb200_mod):
    from advb200 import _lib
    assert _lib.LIB_PATH.is_file()
    before = _lib.launch_count()
    advb200_mod._defense_quantization(torch.rand(1, 3, 8, 8, device="cuda"))
    assert _lib.launch_count() == before + 1


@pytest.mark.parametrize("shape", [(1, 3, 224, 224), (5, 3, 224, 224), (2, 3, 7, 9), (1, 1, 1, 1), (0, 3, 4, 4)])
def test_linf_step_bit_exact(advb200_mod, dev, shape):
    g = torch.Generator().manual_seed(sum(shape))
    x0 = torch.rand(shape, generator=g)
    eps, alpha = 8 / 255, 2 / 255
    x = (x0 + (torch.rand(shape, generator=g) * 2 - 1) * eps).clamp(0, 1)
    gr = torch.randn(shape, generator=g)
    if gr.numel() > 4:
        gr.view(-1)[:3] = torch.tensor([0.0, -0.0, float("nan")])
    want = attacks_ref.linf_update(x, x0, gr, alpha, eps, project=True)
    got = x.to(dev).clone()
    advb200_mod.attack._linf_step(got, x0.to(dev), gr.to(dev), alpha, eps, True)
    assert torch.equal(got.cpu(), want)
    # the constraint the reference guarantees: inside the fp32 eps-ball and inside [0,1]
    assert bool(((got.cpu() <= x0 + eps) & (got.cpu() >= x0 - eps) & (got.cpu() >= 0) & (got.cpu() <= 1)).all())
    want = attacks_ref.linf_update(x, None, gr, eps, eps, project=False)
    got = x.to(dev).clone()
    advb200_mod.attack._linf_step(got, None, gr.to(dev), eps, eps, False)
    assert torch.equal(got.cpu(), want)


def test_linf_step_unaligned_views(advb200_mod, dev):
    g = torch.Generator().manual_seed(3)
    buf = torch.rand(1 + 2 * 3 * 5 * 7, generator=g)
    x0 = buf[1:].view(2, 3, 5, 7)                     # 4-byte aligned only
    gr = torch.randn(2, 3, 5, 7, generator=g)
    want = attacks_ref.linf_update(x0, x0, gr, 0.01, 0.03, project=True)
    dbuf = buf.to(dev)
    got = dbuf[1:].view(2, 3, 5, 7).clone()
    xo = dbuf[1:].view(2, 3, 5, 7)
    from advb200 import _lib
    rc = _lib.load().advb200_linf_step_f32(_lib.ptr(got), _lib.ptr(xo), _lib.ptr(gr.to(dev)), 0.01, 0.03,
                                            got.numel(), 1, _lib.stream_ptr())
    assert rc == 0
    assert torch.equal(got.cpu(), want)


def test_defenses_bit_exact_vs_golden_and_oracle(advb200_mod, dev, golden_dir):
    g = np.load(golden_dir / "defense_small.npz")
    x = torch.from_numpy(g["x"]).to(dev)
    assert np.array_equal(advb200_mod._defense_smoothing(x).cpu().numpy(), g["smooth"])
    assert np.array_equal(advb200_mod._defense_quantization(x, 16).cpu().numpy(), g["quant16"])
    assert np.array_equal(advb200_mod._defense_quantization(x, 4).cpu().numpy(), g["quant4"])
    assert np.array_equal(advb200_mod.defend_input(x).cpu().numpy(), g["defend"])
    # full-size seeded batch against the oracle + size-independent properties
    xb = torch.rand(16, 3, 224, 224, generator=torch.Generator().manual_seed(5)) * 1.2 - 0.1
    d = advb200_mod.defend_input(xb.to(dev))
    assert np.array_equal(d.cpu().numpy(), defenses_ref.defend_input(xb.numpy()))
    q = advb200_mod._defense_quantization(xb.to(dev))
    assert torch.equal(advb200_mod._defense_quantization(q), q)                       # idempotent
    assert int(torch.unique(q).numel()) <= 16 and float(q.min()) >= 0 and float(q.max()) <= 1
    assert np.array_equal(advb200_mod._defense_smoothing(xb.to(dev)).cpu().numpy(), defenses_ref.smooth3x3(xb.numpy()))
    ones = torch.ones(1, 1, 224, 224, device=dev)
    s = advb200_mod._defense_smoothing(ones)
    assert float(s[0, 0, 5, 5]) == 1.0 and float(s[0, 0, 0, 0]) == np.float32(4.0) / np.float32(9.0)


def test_median_smoothing_exact_vs_oracle(advb200_mod, dev):
    """Extra defense (3x3 median, zero pad): a selection, so bit-exact against the oracle; ragged widths take the
    scalar path, multiples of four the 128-bit one; plus the properties a median has at any size."""
    gen = torch.Generator().manual_seed(21)
    for shape in ((16, 3, 224, 224), (2, 3, 17, 23), (1, 1, 1, 1), (1, 2, 1, 9), (1, 1, 6, 2), (3, 1, 5, 8)):
        xb = torch.rand(*shape, generator=gen) * 1.2 - 0.1
        got = advb200_mod._defense_median_smoothing(xb.to(dev)).cpu().numpy()
        assert np.array_equal(got, defenses_ref.median3x3(xb.numpy())), shape
    # a misaligned view (plane base not 16-byte aligned) must take the scalar path and still agree
    buf = torch.rand(1 + 2 * 3 * 8 * 8, generator=gen)
    xv = buf[1:].view(2, 3, 8, 8)
    dv = buf.to(dev)[1:].view(2, 3, 8, 8)
    from advb200 import _lib
    out = torch.empty(2, 3, 8, 8, device=dev)
    assert _lib.load().advb200_median3x3_f32(_lib.ptr(dv), _lib.ptr(out), 6, 8, 8, _lib.stream_ptr()) == 0
    assert np.array_equal(out.cpu().numpy(), defenses_ref.median3x3(xv.numpy()))
    ones = torch.ones(1, 1, 224, 224, device=dev)
    m = advb200_mod._defense_median_smoothing(ones)
    assert float(m[0, 0, 5, 5]) == 1.0 and float(m[0, 0, 0, 0]) == 0.0 and float(m[0, 0, 0, 5]) == 1.0
    big = torch.rand(64, 3, 224, 224, device=dev)
    mb = advb200_mod._defense_median_smoothing(big)
    assert float(mb.min()) >= 0.0 and float(mb.max()) <= float(big.max())
    assert _lib.load().advb200_median3x3_f32(_lib.ptr(big), _lib.ptr(big), 192, 224, 224, _lib.stream_ptr()) != 0
    with pytest.raises(ValueError):
        advb200_mod._defense_median_smoothing(torch.rand(3, 8, 8, device=dev))


def test_divide_by_nine_shortcut_is_exact_for_every_float(advb200_mod):
    """The smoothing kernels divide by 9 with one multiply and two FMAs instead of the generic division sequence;
    the device checks it against __fdiv_rn for all 2^32 bit patterns."""
    import ctypes
    from advb200 import _lib
    bad, first = ctypes.c_uint64(123), ctypes.c_uint32(0)
    _lib.check(_lib.load().advb200_test_div9(ctypes.byref(bad), ctypes.byref(first)), "advb200_test_div9")
    assert bad.value == 0, f"{bad.value} mismatches, first at bits 0x{first.value:08x}"


def test_defense_edge_shapes(advb200_mod, dev):
    for shape in [(1, 1, 1, 1), (2, 3, 1, 5), (1, 2, 3, 1), (0, 3, 8, 8), (1, 3, 33, 17)]:
        x = torch.rand(shape, generator=torch.Generator().manual_seed(1)) * 1.4 - 0.2
        got = advb200_mod.defend_input(x.to(dev)).cpu().numpy()
        assert np.array_equal(got, defenses_ref.defend_input(x.numpy())) if x.numel() else got.size == 0


# ------------------------------------------------------------------ model: logits / gradients
@pytest.mark.parametrize("precision,logit_tol,grad_tol,sign_min", [("fp32", 1e-4, 2e-2, 0.998), ("bf16", 1e-2, 0.6, 0.85)])
def test_logits_and_input_gradient_vs_reference_golden(native, dev, golden_dir, golden_meta, precision, logit_tol,
                                                       grad_tol, sign_min):
    g = np.load(golden_dir / "resnet_n2.npz")
    mean, std = weights.mean_std()
    x = weights.synthetic_images(4)[:2].to(dev)
    y = torch.from_numpy(g["y"]).to(dev)
    out = native[precision].input_grad(x, mean, std, y, loss="ce", want_logits=True, want_preds=True, want_loss=True)
    ref_logits, ref_grad = torch.from_numpy(g["logits"]), torch.from_numpy(g["grad"])
    assert rel(out["logits"], ref_logits) < logit_tol
    assert torch.equal(out["preds"].cpu(), y.cpu())
    ce = torch.nn.functional.cross_entropy(ref_logits, y.cpu(), reduction="none")
    np.testing.assert_allclose(out["loss"].cpu().numpy(), ce.numpy(), rtol=0.5 if precision == "bf16" else 1e-3)
    e = rel(out["grad"], ref_grad)
    mask = ref_grad.abs() > 1e-6 * ref_grad.abs().max()
    agree = float((out["grad"].cpu().sign() == ref_grad.sign())[mask].float().mean())
    print(f"[{precision}] grad rel-L2 vs reference fp32 {e:.3e}; vs fp64 {rel(out['grad'], torch.from_numpy(g['grad64'])):.3e}; "
          f"reference's own fp32-vs-fp64 {golden_meta['ref_fp32_vs_fp64']['grad_rel']:.3e}; sign agreement {agree:.5f}")
    assert e < grad_tol and agree > sign_min
    assert not bool(torch.isnan(out["grad"]).any())


def test_fp32_mode_is_as_close_to_fp64_as_the_reference_is(native, dev, golden_dir, golden_meta):
    g = np.load(golden_dir / "resnet_n2.npz")
    mean, std = weights.mean_std()
    x = weights.synthetic_images(4)[:2].to(dev)
    out = native["fp32"].input_grad(x, mean, std, torch.from_numpy(g["y"]).to(dev), loss="ce", want_logits=True)
    assert rel(out["logits"], torch.from_numpy(g["logits64"])) < 1e-5
    ours = rel(out["grad"], torch.from_numpy(g["grad64"]))
    assert ours < 3 * golden_meta["ref_fp32_vs_fp64"]["grad_rel"]


def test_batch_of_one_and_odd_batches_match_batched(native, dev):
    mean, std = weights.mean_std()
    xs = weights.synthetic_images(5).to(dev)
    nm = native["fp32"]
    full = nm.forward(xs, mean, std)["logits"]
    one = torch.cat([nm.forward(xs[i:i + 1], mean, std)["logits"] for i in range(5)])
    assert rel(one, full) < 1e-5
    y = full.argmax(1)
    gfull = nm.input_grad(xs, mean, std, y, inv_n_global=1.0)["grad"]
    g3 = nm.input_grad(xs[3:4], mean, std, y[3:4], inv_n_global=1.0)["grad"]
    assert rel(g3, gfull[3:4]) < 1e-3
    fb = native["bf16"].forward(xs, mean, std)["logits"]
    ob = torch.cat([native["bf16"].forward(xs[i:i + 1], mean, std)["logits"] for i in range(5)])
    assert torch.equal(fb, ob)          # tile contents do not depend on the batch they ride in


def test_cw_margin_gradient_vs_torch_autograd(native, torch_model, dev):
    mean, std = weights.mean_std()
    x = weights.synthetic_images(2).to(dev)
    y = torch.tensor([713, 5], device=dev)
    for targeted in (False, True):
        xi = x.clone().requires_grad_(True)
        logits = torch_model(attacks_ref.normalize(xi, mean.to(dev), std.to(dev)))
        onehot = torch.nn.functional.one_hot(y, 1000).float()
        real = (logits * onehot).sum(1)
        other = (logits - 1e4 * onehot).amax(1)
        f = torch.clamp(other - real + 0.5, min=0) if targeted else torch.clamp(real - other + 0.5, min=0)
        (gref,) = torch.autograd.grad((2.0 * f).sum(), xi)
        out = native["fp32"].input_grad(x, mean, std, y, loss="cw_targeted" if targeted else "cw", c=2.0, kappa=0.5,
                                        want_loss=True, want_success=True)
        np.testing.assert_allclose(out["loss"].cpu().numpy(), f.detach().cpu().numpy(), rtol=1e-4, atol=1e-4)
        if float(gref.abs().max()) > 0:
            assert rel(out["grad"], gref) < 5e-2
        else:
            assert float(out["grad"].abs().max()) == 0.0
        succ = (logits.argmax(1) == y) if targeted else (logits.argmax(1) != y)
        assert torch.equal(out["success"].bool().cpu(), succ.cpu())


# ------------------------------------------------------------------ attacks end to end
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_fgsm_pgd_constraints_and_success_vs_golden(advb200_mod, native, dev, golden_dir, precision):
    g = np.load(golden_dir / "attacks_n2.npz")
    r = np.load(golden_dir / "resnet_n2.npz")
    mean, std = weights.mean_std()
    x = weights.synthetic_images(4)[:2].to(dev)
    y = torch.from_numpy(r["y"]).to(dev)
    eps, alpha = 8 / 255, 2 / 255
    nm = native[precision]
    xf = advb200_mod.fgsm_attack(nm, x, y, eps=eps, mean=mean, std=std)
    assert float((xf - x).abs().max()) <= eps + 1e-7 and float(xf.min()) >= 0 and float(xf.max()) <= 1
    # FGSM output is x + eps*sign(g): compare the sign pattern with the reference gradient
    ref_sign = torch.from_numpy(r["grad"]).sign()
    mask = torch.from_numpy(np.abs(r["grad"]) > 1e-6 * np.abs(r["grad"]).max())
    interior = ((x.cpu() + eps <= 1) & (x.cpu() - eps >= 0)) & mask
    got_sign = ((xf - x).cpu() / eps).round()
    agree = float((got_sign == ref_sign)[interior].float().mean())
    assert agree > (0.998 if precision == "fp32" else 0.85)
    pf = nm.forward(xf, mean, std, want_preds=True)["preds"].cpu()
    assert torch.equal(pf != y.cpu(), torch.from_numpy(g["fgsm_logits"]).argmax(1) != y.cpu())   # same success flags
    torch.manual_seed(7)
    xp = advb200_mod.pgd_linf_attack(nm, x, y, eps=eps, alpha=alpha, steps=10, mean=mean, std=std)
    hi, lo = x + eps, x - eps                     # fp32 arithmetic, as the reference evaluates it
    assert bool(((xp <= hi) & (xp >= lo) & (xp >= 0) & (xp <= 1)).all())
    pp = nm.forward(xp, mean, std, want_preds=True)["preds"].cpu()
    assert torch.equal(pp != y.cpu(), torch.from_numpy(g["pgd_logits"]).argmax(1) != y.cpu())
    x_same = x.clone()
    advb200_mod.pgd_linf_attack(nm, x, y, eps=eps, alpha=alpha, steps=1, mean=mean, std=std, random_start=False)
    assert torch.equal(x, x_same)                 # inputs are never mutated


def test_pgd_tracks_the_oracle_step_by_step_fp32(advb200_mod, native, dev):
    """Same start noise, fp32 mode: after each step the iterate equals the oracle's on all pixels
    whose gradient sign agrees; the disagreement stays at the reference's own noise level."""
    mean, std = weights.mean_std()
    x = weights.synthetic_images(1)
    fm = resnet50_ref.FunctionalResNet50(weights.seeded_resnet50(0).state_dict())
    y = fm(attacks_ref.normalize(x, mean, std)).argmax(1)
    eps, alpha = 8 / 255, 2 / 255
    noise = weights.pgd_noise(x.shape, eps, 3)
    xa = (x + noise).clamp(0, 1).to(dev)          # both sides start from the oracle's noise tensor
    xo = x.to(dev)
    fracs = []
    for steps in (1, 2):
        want = attacks_ref.pgd_linf(fm, x, y, eps=eps, alpha=alpha, steps=steps, mean=mean, std=std, noise=noise)
        gq = native["fp32"].input_grad(xa, mean, std, y.to(dev), loss="ce")["grad"]
        advb200_mod.attack._linf_step(xa, xo, gq, alpha, eps, True)
        fracs.append(float((xa.cpu() == want).float().mean()))
    print(f"PGD pixels identical to the oracle after 1 / 2 steps: {fracs[0]:.5f} / {fracs[1]:.5f}")
    # step 1 differs only where the gradient sign differs (the reference's own fp32 noise: ~5e-4 of
    # pixels); the model is chaotic in its input, so the disagreement compounds from step 2 on
    assert fracs[0] > 0.998
    assert fracs[1] > 0.9


def test_cw_l2_matches_reference_golden_fp32(advb200_mod, native, dev, golden_dir):
    g = np.load(golden_dir / "attacks_n2.npz")
    r = np.load(golden_dir / "resnet_n2.npz")
    mean, std = weights.mean_std()
    x = weights.synthetic_images(4)[:2].to(dev)
    y = torch.from_numpy(r["y"]).to(dev)
    nm = native["fp32"]
    res, l2 = advb200_mod.cw_l2_attack(nm, x, y, mean=mean, std=std, c=1.0, kappa=0.0, steps=20, lr=0.01, return_l2=True)
    assert np.array_equal(res.success.cpu().numpy(), g["cwu_success"])
    got_l2 = (res.x_adv - x.clamp(0, 1)).flatten(1).pow(2).sum(1).cpu().numpy()
    np.testing.assert_allclose(got_l2, g["cwu_l2"], rtol=2e-2)
    ok = g["cwu_success"].astype(bool)
    np.testing.assert_allclose(l2.cpu().numpy()[ok], got_l2[ok], rtol=1e-4)
    res = advb200_mod.cw_l2_attack(nm, x, y, mean=mean, std=std, c=1.0, kappa=0.0, steps=20, lr=0.01, targeted=True,
                                   y_target=torch.full((2,), 805, device=dev))
    assert np.array_equal(res.success.cpu().numpy(), g["cwt_success"])
    got_l2 = (res.x_adv - x.clamp(0, 1)).flatten(1).pow(2).sum(1).cpu().numpy()
    np.testing.assert_allclose(got_l2, g["cwt_l2"], rtol=2e-2)
    assert float(res.x_adv.min()) >= 0 and float(res.x_adv.max()) <= 1


def test_cw_l2_bf16_success_flags(advb200_mod, native, dev, golden_dir):
    g = np.load(golden_dir / "attacks_n2.npz")
    r = np.load(golden_dir / "resnet_n2.npz")
    mean, std = weights.mean_std()
    x = weights.synthetic_images(4)[:2].to(dev)
    y = torch.from_numpy(r["y"]).to(dev)
    res = advb200_mod.cw_l2_attack(native["bf16"], x, y, mean=mean, std=std, steps=20, lr=0.01)
    assert np.array_equal(res.success.cpu().numpy(), g["cwu_success"])


# ------------------------------------------------------------------ detector / evaluation
def test_detector_scores_and_threshold_vs_golden(advb200_mod, native, dev, golden_meta):
    mean, std = weights.mean_std()
    xs = weights.synthetic_images(4).to(dev)
    det = golden_meta["detector"]
    want = np.array(det["scores"], dtype=np.float32)
    s32 = advb200_mod._layer3_features(native["fp32"], xs, mean, std).cpu().numpy()
    np.testing.assert_allclose(s32, want, rtol=2e-5)
    sbf = advb200_mod._layer3_features(native["bf16"], xs, mean, std).cpu().numpy()
    np.testing.assert_allclose(sbf, want, rtol=5e-3)
    thr = advb200_mod.calibrate_feature_threshold(native["fp32"], xs, mean=mean, std=std, n=100, quantile=0.95)
    assert abs(thr - det["threshold"]) < 1e-3
    flags = advb200_mod.is_adversarial_by_feature(native["fp32"], xs, mean, std, det["threshold"]).cpu().numpy()
    assert np.array_equal(flags, want > det["threshold"])


def test_evaluate_defenses_matches_reference_batch1_dicts(advb200_mod, native, dev, golden_meta):
    """fp32 validation mode reproduces the reference's per-image booleans (batch-1 loop)."""
    mean, std = weights.mean_std()
    xs = weights.synthetic_images(4).to(dev)
    det = golden_meta["detector"]
    nm = native["fp32"]
    for key, rows in det["evaluate_defenses"].items():
        name, e = key.split("_")
        for i, want in enumerate(rows):
            x = xs[i:i + 1]
            y = nm.forward(x, mean, std, want_preds=True)["preds"]
            torch.manual_seed(100 + i)
            got = advb200_mod.evaluate_defenses(nm, x, y, mean, std, name, int(e) / 255, 2 / 255, 10, 1.0, 0.0, 20,
                                                0.01, det["threshold"])
            got = {k: int(v) for k, v in got.items()}
            # the two detector flags sit within 1e-3 of the threshold for some images:
            # compare everything, and report rather than hide a detector flip
            for k in ("clean_correct", "attack_success", "defense_preproc_success"):
                assert got[k] == want[k], (key, i, k)
            assert got["detector_flags_clean"] == want["detector_flags_clean"], (key, i)
            assert got["detector_flags_adv"] == want["detector_flags_adv"], (key, i)
            assert got["detector_attack_success"] == want["detector_attack_success"], (key, i)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_batched_evaluator_equals_batch1_loop(advb200_mod, native, dev, golden_meta, precision):
    mean, std = weights.mean_std()
    xs = weights.synthetic_images(6).to(dev)
    nm = native[precision]
    thr = golden_meta["detector"]["threshold"]
    y = nm.forward(xs, mean, std, want_preds=True)["preds"]
    counters, flags = advb200_mod.evaluate_defenses_batched(nm, xs, y, mean, std, "fgsm", 4 / 255, 2 / 255, 10, 1.0,
                                                            0.0, 20, 0.01, thr, return_flags=True)
    loop = torch.zeros(7, dtype=torch.int64)
    for i in range(6):
        c1, f1 = advb200_mod.evaluate_defenses_batched(nm, xs[i:i + 1], y[i:i + 1], mean, std, "fgsm", 4 / 255, 2 / 255,
                                                       10, 1.0, 0.0, 20, 0.01, thr, return_flags=True)
        loop += c1.cpu()
        assert torch.equal(f1[0].cpu(), flags[i].cpu()), i
    assert torch.equal(loop, counters.cpu())
    assert int(counters[6]) == 6 and int(counters[0]) == 6


def test_config1_top5_lines_vs_reference_cli(advb200_mod, native, dev, golden_dir, golden_meta):
    """Config 1: the reference CLI's 2x5 printed lines on picture/example.jpg (random-init weights)."""
    import re
    mean, std = weights.mean_std()
    x = (torch.from_numpy(np.load(golden_dir / "example_224_u8.npy")).float() / 255).unsqueeze(0).to(dev)
    nm = native["fp32"]
    lines = golden_meta["cli_config1_stdout"]
    parse = lambda ls: [(int(re.search(r"class (\d+)", l).group(1)), float(l.rsplit("= ", 1)[1])) for l in ls]
    clean_want, adv_want = parse(lines[2:7]), parse(lines[8:13])
    logits = nm.forward(x, mean, std)["logits"]
    p, idx = torch.softmax(logits, 1).topk(5, 1)
    assert idx[0].tolist() == [c for c, _ in clean_want]
    np.testing.assert_allclose(p[0].cpu().numpy(), [q for _, q in clean_want], atol=1e-4)
    xa = advb200_mod.fgsm_attack(nm, x, logits.argmax(1), eps=0.031372549, mean=mean, std=std)
    p, idx = torch.softmax(nm.forward(xa, mean, std)["logits"], 1).topk(5, 1)
    assert idx[0].tolist() == [c for c, _ in adv_want]
    np.testing.assert_allclose(p[0].cpu().numpy(), [q for _, q in adv_want], atol=2e-3)


# ------------------------------------------------------------------ full-size properties (config 2)
def test_pgd10_batch256_bf16_properties(advb200_mod, native, dev):
    mean, std = weights.mean_std()
    x = weights.synthetic_images(256).to(dev)
    nm = native["bf16"]
    y = nm.forward(x, mean, std, want_preds=True)["preds"]
    eps, alpha = 8 / 255, 2 / 255
    torch.manual_seed(11)
    xa = advb200_mod.pgd_linf_attack(nm, x, y, eps=eps, alpha=alpha, steps=10, mean=mean, std=std)
    assert bool(((xa <= x + eps) & (xa >= x - eps) & (xa >= 0) & (xa <= 1)).all())
    pa = nm.forward(xa, mean, std, want_preds=True)["preds"]
    assert float((pa != y).float().mean()) == 1.0          # reference: PGD-10 success 1.000 on this model
    # every pixel moved by a multiple of alpha steps inside the ball unless clamped: |delta| <= eps
    assert float((xa - x).abs().max()) <= eps + 1e-7
    # sharded == unsharded: a rank that owns images 64..127 computes the same iterate for them
    torch.manual_seed(11)
    noise = torch.empty_like(x).uniform_(-eps, eps)
    lo, hi = 64, 128
    xs = (x[lo:hi] + noise[lo:hi]).clamp(0, 1)
    g_shard = nm.input_grad(xs, mean, std, y[lo:hi], inv_n_global=1 / 256)["grad"]
    g_full = nm.input_grad((x + noise).clamp(0, 1), mean, std, y, inv_n_global=1 / 256)["grad"][lo:hi]
    assert torch.equal(g_shard, g_full)


@pytest.mark.gpu
def test_host_batch_pipeline_equals_direct_calls(advb200_mod, native, dev):
    """The overlapped host-buffer front end returns exactly what direct calls on device tensors return."""
    import functools
    from oracle import weights
    nm = native["bf16"]
    mean, std = weights.mean_std()
    batches = [weights.synthetic_images(3, seed=50 + i).pin_memory() for i in range(4)]
    y = torch.zeros(3, dtype=torch.int64, device=dev)
    fn = functools.partial(advb200_mod.fgsm_attack, nm, y_true=y, eps=8 / 255, mean=mean, std=std)
    pipe = advb200_mod.HostBatchPipeline(fn, dev)
    got = [r.clone() for r in pipe.run(batches)]
    assert len(got) == 4
    for b, r in zip(batches, got):
        want = fn(b.to(dev)).cpu()
        assert torch.equal(r, want)
    # caller-supplied result buffers, two used in turn: batch i-2 is handed out before batch i overwrites its buffer
    import itertools
    outs = [torch.empty(3, 3, 224, 224).pin_memory() for _ in range(2)]
    more = batches + [weights.synthetic_images(3, seed=60).pin_memory()]
    got2 = [r.clone() for r in pipe.run(more, itertools.cycle(outs))]
    assert len(got2) == 5
    for b, r in zip(more, got2):
        assert torch.equal(r, fn(b.to(dev)).cpu())
    assert [r.clone() for r in pipe.run(batches[:1])][0].equal(got[0])          # a single batch drains correctly
    with pytest.raises(ValueError):
        list(pipe.run([torch.zeros(3, 3, 224, 224)]))


@pytest.mark.gpu
@pytest.mark.parametrize("project", [True, False])
def test_fused_attack_step_is_bit_identical_to_gradient_plus_update(advb200_mod, native, dev, project):
    """advb200_attack_step (update fused into the stem input-gradient epilogue) == input_grad + linf step."""
    nm = native["bf16"]
    mean, std = weights.mean_std()
    x0 = weights.synthetic_images(5, seed=77).to(dev)
    y = torch.tensor([1, 713, 5, 713, 999], dtype=torch.int64, device=dev)
    xa = (x0 + 0.01 * torch.randn_like(x0)).clamp(0, 1).contiguous()
    g = nm.input_grad(xa, mean, std, y, loss="ce")["grad"]
    want = xa.clone()
    advb200_mod.attack._linf_step(want, x0 if project else None, g, 2 / 255, 8 / 255, project)
    got = xa.clone()
    nm.attack_step(got, x0 if project else None, mean, std, y, alpha=2 / 255, eps=8 / 255)
    assert torch.equal(got, want)
    assert not torch.equal(got, xa)
    with pytest.raises(RuntimeError):
        native["fp32"].attack_step(got, x0, mean, std, y, alpha=2 / 255, eps=8 / 255)
