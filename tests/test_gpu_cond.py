"""Model-level parity on the CONDITIONED fixture model (oracle/weights.conditioned_resnet50: non-trivial scale and shift in
all 53 BatchNorms, O(1) activations, predictions that differ between images, success rates strictly inside (0,1) at the
small eps points) against vectors the reference itself produced on it (oracle/gen_golden_models.py --model cond).

What can and cannot be asserted (measured in tests/golden/meta_cond.json): the reference's own fp32 input
gradient is 1e-3 away from its fp64 gradient on this model, and its own all-bf16 run is 0.40 away from its fp32 run with
86 % sign agreement — rounding activations to bf16 flips ReLU gates and the network amplifies it.  So: logits are held to
the north-star tolerances (1e-4 fp32 / 1e-2 bf16); the fp32-mode gradient is held to the reference's own fp64 distance;
the bf16 gradient is held to the reference's own bf16 figures and printed beside them, and its kernels are pinned
launch by launch to one ulp in tests/test_gpu_trace.py instead."""
import json

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import weights  # noqa: E402


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def meta(golden_dir):
    return json.loads((golden_dir / "meta_cond.json").read_text())


@pytest.fixture(scope="module")
def native(dev):
    import advb200
    model = weights.conditioned_resnet50(0).to(dev)
    return {p: advb200.NativeResNet50(model, precision=p) for p in ("fp32", "bf16")}


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / b.norm())


def images(n=8):
    return weights.mixed_images(8, seed=77)[:n]


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_logits_and_input_gradient_vs_reference_golden(native, dev, golden_dir, meta, precision):
    g = np.load(golden_dir / "cond_resnet.npz")
    mean, std = weights.mean_std()
    x = images(4).to(dev)
    y = torch.from_numpy(g["y"]).to(dev)
    out = native[precision].input_grad(x, mean, std, y, loss="ce", want_logits=True, want_preds=True, want_loss=True)
    ref_logits, ref_grad, g64 = torch.from_numpy(g["logits"]), torch.from_numpy(g["grad"]), torch.from_numpy(g["grad64"])
    lrel = rel(out["logits"], ref_logits)
    e32, e64 = rel(out["grad"], ref_grad), rel(out["grad"], g64)
    mask = g64.abs() > 1e-6 * g64.abs().max()
    agree = float((out["grad"].cpu().sign() == g64.sign())[mask].float().mean())
    own32, ownbf = meta["ref_fp32_vs_fp64"], meta["ref_bf16_vs_fp32"]
    print(f"[cond/{precision}] logits rel-L2 {lrel:.2e}; grad rel-L2 vs reference fp32 {e32:.3e}, vs fp64 {e64:.3e}, sign agreement "
          f"{agree:.5f} | the reference's own: fp32 vs fp64 {own32['grad_rel']:.3e} ({own32['sign_mismatches']} sign flips of "
          f"{own32['masked']}), all-bf16 vs fp32 {ownbf['grad_rel']:.3f} / {ownbf['sign_agreement']:.4f} (logits {ownbf['logits_rel']:.2e})")
    assert not bool(torch.isnan(out["grad"]).any())
    ce = torch.nn.functional.cross_entropy(ref_logits, y.cpu(), reduction="none")
    if precision == "fp32":
        assert lrel < 1e-4                                              # north star, fp32 validation mode
        assert torch.equal(out["preds"].cpu(), y.cpu())
        np.testing.assert_allclose(out["loss"].cpu().numpy(), ce.numpy(), rtol=1e-4)
        assert e64 < 3 * own32["grad_rel"]                              # as close to fp64 as the reference itself is
        assert agree > 1 - 5 * own32["sign_mismatches"] / own32["masked"]
    else:
        assert lrel < 1e-2                                              # north star, bf16
        np.testing.assert_allclose(out["loss"].cpu().numpy(), ce.numpy(), rtol=2e-2)
        assert e32 < 1.25 * ownbf["grad_rel"] and agree > ownbf["sign_agreement"] - 0.03


def test_bf16_gradient_beside_the_bf16_emulator(native, dev, golden_dir):
    """Why the whole-network bf16 gradient cannot be pinned tighter than bf16 noise: the bf16-emulating oracle
    (oracle/resnet50_bf16emu.py: the reference network with a bf16 rounding at every point the product path stores bf16,
    float64 sums) is itself as far from the native bf16 gradient as two such emulations are from each other when they
    differ only in accumulation precision.  Reported; the per-launch comparison (tests/test_gpu_trace.py) is the pin."""
    from oracle import resnet50_bf16emu as emu
    g = np.load(golden_dir / "cond_resnet.npz")
    mean, std = weights.mean_std()
    x = images(2).to(dev)
    y = torch.from_numpy(g["y"][:2]).to(dev)
    sd = weights.conditioned_resnet50(0).state_dict()
    e64 = emu.Bf16EmuResNet50(sd, emulate=True, dtype=torch.float64, device=dev)
    e32 = emu.Bf16EmuResNet50(sd, emulate=True, dtype=torch.float32, device=dev)
    g64, l64 = e64.input_grad(x, y, mean, std, inv_n=1 / 4)
    g32, l32 = e32.input_grad(x, y, mean, std, inv_n=1 / 4)
    out = native["bf16"].input_grad(x, mean, std, y, loss="ce", want_logits=True, inv_n_global=1 / 4)
    ref_grad = torch.from_numpy(g["grad"][:2])
    r_nat_emu, r_emu_emu = rel(out["grad"], g64), rel(g32, g64)
    print(f"[cond/bf16] grad rel-L2: native vs bf16 emulator (fp64 sums) {r_nat_emu:.3f}; emulator with fp32 sums vs fp64 sums {r_emu_emu:.3f}; "
          f"native vs fp32 reference {rel(out['grad'], ref_grad):.3f}; emulator vs fp32 reference {rel(g64, ref_grad):.3f}; "
          f"logits native vs emulator {rel(out['logits'], l64):.2e}")
    assert rel(out["logits"], l64) < 1e-2
    assert r_nat_emu < 2.5 * max(r_emu_emu, 0.1)            # no further from the emulator than the emulator is from itself


def test_fgsm_pgd_outcomes_vs_reference_golden_fp32(native, dev, golden_dir):
    """FGSM pixels identical to the reference's output wherever the gradient sign agrees; success flags of FGSM and PGD-10
    equal the reference's, at eps 8/255 (all succeed) and at the small eps where they do not."""
    import advb200
    a = np.load(golden_dir / "cond_attacks.npz")
    r = np.load(golden_dir / "cond_resnet.npz")
    mean, std = weights.mean_std()
    x = images(2).to(dev)
    y = torch.from_numpy(r["y"][:2]).to(dev)
    nm = native["fp32"]
    seen = set()
    for tag in ("8", "small"):
        eps = float(a["eps_" + tag])
        xf = advb200.fgsm_attack(nm, x, y, eps=eps, mean=mean, std=std)
        same = float((xf.cpu() == torch.from_numpy(a["fgsm_x_" + tag])).float().mean())
        assert same > 0.998, (tag, same)
        pf = nm.forward(xf, mean, std, want_preds=True)["preds"].cpu()
        want_f = torch.from_numpy(a["fgsm_logits_" + tag]).argmax(1) != y.cpu()
        assert torch.equal(pf != y.cpu(), want_f), tag
        noise = weights.pgd_noise(x.shape, eps, 7)                       # what attack.py:51 drew after manual_seed(7)
        xp = advb200.pgd_linf_attack(nm, x, y, eps=eps, alpha=2 / 255, steps=10, mean=mean, std=std, noise=noise.to(dev))
        assert bool(((xp <= x + eps) & (xp >= x - eps) & (xp >= 0) & (xp <= 1)).all())
        pp = nm.forward(xp, mean, std, want_preds=True)["preds"].cpu()
        want_p = torch.from_numpy(a["pgd_logits_" + tag]).argmax(1) != y.cpu()
        assert torch.equal(pp != y.cpu(), want_p), tag
        seen |= set(want_f.tolist()) | set(want_p.tolist())
        print(f"[cond] eps={eps * 255:g}/255: FGSM pixels identical to the reference {same:.5f}; success FGSM {want_f.tolist()} "
              f"PGD {want_p.tolist()} (equal to the reference)")
    assert seen == {True, False}                                         # the fixture really has both outcomes


def test_cw_l2_vs_reference_golden_fp32(native, dev, golden_dir):
    import advb200
    a = np.load(golden_dir / "cond_attacks.npz")
    r = np.load(golden_dir / "cond_resnet.npz")
    mean, std = weights.mean_std()
    x = images(2).to(dev)
    y = torch.from_numpy(r["y"][:2]).to(dev)
    nm = native["fp32"]
    for tag, kw in (("cwu", {}), ("cwt", {"targeted": True, "y_target": torch.full((2,), 805, device=dev)})):
        res = advb200.cw_l2_attack(nm, x, y, mean=mean, std=std, c=1.0, kappa=0.0, steps=20, lr=0.01, **kw)
        assert np.array_equal(res.success.cpu().numpy(), a[tag + "_success"])
        l2 = (res.x_adv - x.clamp(0, 1)).flatten(1).pow(2).sum(1).cpu().numpy()
        err = float(np.abs(l2 / a[tag + "_l2"] - 1).max())
        pred = nm.forward(res.x_adv, mean, std, want_preds=True)["preds"].cpu().numpy()
        print(f"[cond] CW-L2 20 steps {tag}: success {res.success.tolist()} (= reference), l2 {l2.tolist()} vs {a[tag + '_l2'].tolist()} "
              f"(max rel err {err:.2e}); final predictions {pred.tolist()} vs {a[tag + '_logits'].argmax(1).tolist()}")
        # the reference's own spread of this number (8 threads vs 1 thread, fp32 vs fp64) is 2.5e-3 on this model
        # (oracle/probe_cw_noise.py): Adam's unit-magnitude steps turn a gradient sign flip into a 2*lr move
        assert err < 1e-2
        assert np.array_equal(pred, a[tag + "_logits"].argmax(1))
    bf = advb200.cw_l2_attack(native["bf16"], x, y, mean=mean, std=std, c=1.0, kappa=0.0, steps=20, lr=0.01)
    assert np.array_equal(bf.success.cpu().numpy(), a["cwu_success"])
    l2b = (bf.x_adv - x.clamp(0, 1)).flatten(1).pow(2).sum(1).cpu().numpy()
    print(f"[cond] CW-L2 20 steps bf16: l2 {l2b.tolist()} vs reference {a['cwu_l2'].tolist()}")
    np.testing.assert_allclose(l2b, a["cwu_l2"], rtol=0.1)


def test_detector_and_evaluate_defenses_vs_reference_golden(native, dev, meta):
    """Per-image detector scores, and the six booleans of evaluate_defenses for 2 images x 8 (attack, eps) settings, equal
    to the reference's batch-1 loop (fp32 mode, the reference's PGD start noise injected)."""
    import advb200
    mean, std = weights.mean_std()
    xs = images(8).to(dev)
    det = meta["detector"]
    want = np.array(det["scores"], dtype=np.float32)
    s32 = advb200._layer3_features(native["fp32"], xs, mean, std).cpu().numpy()
    np.testing.assert_allclose(s32, want, rtol=2e-5)
    sbf = advb200._layer3_features(native["bf16"], xs, mean, std).cpu().numpy()
    np.testing.assert_allclose(sbf, want, rtol=5e-3)
    print(f"[cond] detector scores: fp32 max rel err {np.abs(s32 / want - 1).max():.1e}, bf16 {np.abs(sbf / want - 1).max():.1e}; "
          f"threshold {det['threshold']:.5f}, closest clean score margin {np.abs(want - det['threshold']).min():.1e}")
    nm = native["fp32"]
    thr = det["threshold"]
    keys = ("clean_correct", "attack_success", "defense_preproc_success", "detector_flags_clean", "detector_flags_adv",
            "detector_attack_success")
    sums = np.zeros(6, dtype=int)
    near_ties = []
    for key, rows in det["evaluate_defenses"].items():
        name, e = key.split("_")
        eps = float(e) / 255
        for i, w in enumerate(rows):
            x = xs[i:i + 1]
            y = nm.forward(x, mean, std, want_preds=True)["preds"]
            noise = weights.pgd_noise(x.shape, eps, 100 + i).to(dev) if name == "pgd" else None
            _, flags = advb200.evaluate_defenses_batched(nm, x, y, mean, std, name, eps, 2 / 255, 10, 1.0, 0.0, 20, 0.01, thr,
                                                         return_flags=True, pgd_noise=noise)
            got = flags[0].cpu().tolist()
            wantl = [w[k] for k in keys]
            if got != wantl:
                # the model's decisions are near-ties (top-1/top-2 logit gaps of 1e-3..1e-1): a prediction may differ only
                # where the REFERENCE's own decision gap for that forward was below 0.02, and it is reported
                gap_clean, gap_adv, gap_def = det["decision_gaps"][key][i]
                col_gap = {0: gap_clean, 1: gap_adv, 2: gap_def, 5: gap_adv}
                for col in range(6):
                    if got[col] != wantl[col]:
                        assert col in col_gap and col_gap[col] < 0.02, (key, i, keys[col], got, wantl, det["decision_gaps"][key][i])
                        near_ties.append((key, i, keys[col], round(col_gap[col], 4)))
            sums += np.array(got)
    total = 2 * len(det["evaluate_defenses"])
    assert 0 < sums[1] < total and 0 < sums[2] < total and 0 < sums[4] < total      # non-degenerate rates
    assert len(near_ties) <= 2, near_ties
    print(f"[cond] evaluate_defenses: {total} (image, attack, eps) rows; {total * 6 - len(near_ties)} of {total * 6} booleans equal to the "
          f"reference, the rest near-ties of the reference itself (key, image, flag, reference top-2 gap): {near_ties}; sums {dict(zip(keys, sums.tolist()))}")
