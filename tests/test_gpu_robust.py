"""`--model_type robust` (defense_experiments.py:672-686, 764-767): the same ResNet-50 architecture behind a model that
normalises internally, so the attack functions are called with mean = 0 / std = 1.  The RobustBench checkpoint is not
available offline; the CODE PATH is: the normalisation fused into the stem becomes the identity, and every attack /
defense function must take mean (0,0,0) / std (1,1,1) like any other statistics.  Checked against the oracle (the
reference's formulas on the same seeded weights) in the fp32 validation mode, and end to end through the CLI."""
import copy
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import attacks_ref, defenses_ref, resnet50_ref, weights  # noqa: E402

ROOT = Path(__file__).resolve().parent.parent
CLI = ROOT / "image-recognition-adversarial-example-attack_b200" / "cli" / "defense_experiments.py"


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def cond():
    return weights.conditioned_resnet50(0)


def test_identity_normalisation_through_the_attack_functions_fp32(dev, cond):
    """mean = 0 / std = 1 through forward, input gradient, FGSM, PGD and evaluate-style defenses: logits 1e-4 and the
    gradient as close to the oracle as with the ImageNet statistics; the update rules stay bit-exact on the native gradient."""
    import advb200
    mean, std = torch.zeros(3), torch.ones(3)
    x = weights.mixed_images(8, seed=77)[:2]
    fm = resnet50_ref.FunctionalResNet50(cond.state_dict())
    nm = advb200.NativeResNet50(copy.deepcopy(cond).to(dev), precision="fp32")
    with torch.no_grad():
        y = fm(attacks_ref.normalize(x, mean, std)).argmax(1)
    g_ref, logits_ref = attacks_ref.ce_input_grad(fm, x, y, mean, std)
    out = nm.input_grad(x.to(dev), mean, std, y.to(dev), loss="ce", want_logits=True)
    lrel = float((out["logits"].cpu() - logits_ref).norm() / logits_ref.norm())
    grel = float((out["grad"].cpu() - g_ref).norm() / g_ref.norm())
    mask = g_ref.abs() > 1e-6 * g_ref.abs().max()
    agree = float((out["grad"].cpu().sign() == g_ref.sign())[mask].float().mean())
    print(f"[robust path] logits rel {lrel:.2e}, grad rel {grel:.2e}, sign agreement {agree:.5f}")
    assert lrel < 1e-4
    assert grel < 5e-3 and agree > 0.998           # two fp32 evaluations of the same gradient (different summation orders)
    # the statistics really are plumbed through: the ImageNet ones give other logits
    m2, s2 = weights.mean_std()
    other = nm.forward(x.to(dev), m2, s2)["logits"].cpu()
    assert float((other - logits_ref).norm() / logits_ref.norm()) > 1e-2
    # FGSM / PGD: the reference's update applied to the NATIVE gradient, bit for bit; constraints hold
    eps, alpha = 8 / 255, 2 / 255
    xf = advb200.fgsm_attack(nm, x.to(dev), y.to(dev), eps=eps, mean=mean, std=std).cpu()
    want = attacks_ref.linf_update(x, None, out["grad"].cpu(), eps, eps, project=False)
    assert torch.equal(xf, want)
    same = float((xf == attacks_ref.fgsm(fm, x, y, eps=eps, mean=mean, std=std)).float().mean())
    assert same > 0.998, same                      # and the oracle's own FGSM wherever the gradient sign agrees
    noise = weights.pgd_noise(x.shape, eps, 7)
    xp = advb200.pgd_linf_attack(nm, x.to(dev), y.to(dev), eps=eps, alpha=alpha, steps=3, mean=mean, std=std,
                                 noise=noise.to(dev)).cpu()
    assert bool(((xp <= x + eps) & (xp >= x - eps) & (xp >= 0) & (xp <= 1)).all())
    # one step from the same random start: identical wherever the gradient sign agrees (the FGSM bar); over three steps
    # the rare sign flips compound, so the trajectories are compared by distance, not pixel identity
    x1 = advb200.pgd_linf_attack(nm, x.to(dev), y.to(dev), eps=eps, alpha=alpha, steps=1, mean=mean, std=std,
                                 noise=noise.to(dev)).cpu()
    o1 = attacks_ref.pgd_linf(fm, x, y, eps=eps, alpha=alpha, steps=1, mean=mean, std=std, noise=noise)
    assert float((x1 == o1).float().mean()) > 0.998
    xo = attacks_ref.pgd_linf(fm, x, y, eps=eps, alpha=alpha, steps=3, mean=mean, std=std, noise=noise)
    same3 = float((xp == xo).float().mean())
    print(f"[robust path] PGD: 1 step identical {float((x1 == o1).float().mean()):.5f}, 3 steps identical {same3:.4f}")
    assert same3 > 0.9 and float((xp - xo).abs().max()) <= 2 * 3 * alpha + 1e-6
    with torch.no_grad():
        assert torch.equal(fm(attacks_ref.normalize(xp, mean, std)).argmax(1),
                           nm.forward(xp.to(dev), mean, std, want_preds=True)["preds"].cpu())
    # the defenses do not depend on the statistics; the detector score does (it runs the network)
    xd = advb200.defend_input(xp.to(dev)).cpu().numpy()
    assert np.array_equal(xd, defenses_ref.defend_input(xp.numpy()))


def test_identity_normalisation_bf16_logits(dev, cond):
    import advb200
    mean, std = torch.zeros(3), torch.ones(3)
    x = weights.mixed_images(8, seed=77)[:4]
    fm = resnet50_ref.FunctionalResNet50(cond.state_dict())
    with torch.no_grad():
        ref = fm(attacks_ref.normalize(x, mean, std))
    got = advb200.NativeResNet50(copy.deepcopy(cond).to(dev), precision="bf16").forward(x.to(dev), mean, std)["logits"].cpu()
    lrel = float((got - ref).norm() / ref.norm())
    print(f"[robust path] bf16 logits rel {lrel:.2e}")
    assert lrel < 1e-2


def test_cli_model_type_robust(tmp_path):
    """`--model_type robust --weights random` runs the sweep with mean 0 / std 1 and prints the reference's summary lines;
    without `--weights random` it says why the RobustBench checkpoint cannot be used here."""
    cmd = [sys.executable, str(CLI), "--model_type", "robust", "--weights", "random", "--synthetic", "16", "--attacks", "fgsm", "pgd",
           "--eps_list", str(8 / 255), "--steps", "3", "--output_dir", str(tmp_path)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "detector threshold" in r.stdout
    rows = [ln for ln in r.stdout.splitlines() if ln.startswith("counters")]
    assert len(rows) == 2, r.stdout                # one int64[7] counter line per (attack, eps)
    std_cmd = [c for c in cmd if c != "robust"]
    std_cmd[std_cmd.index("--model_type") + 1:std_cmd.index("--model_type") + 1] = ["standard"]
    r2 = subprocess.run(std_cmd, capture_output=True, text=True, timeout=600)
    assert r2.returncode == 0, r2.stderr[-2000:]
    thr = lambda out: [ln for ln in out.splitlines() if "detector threshold" in ln][0]
    assert thr(r.stdout) != thr(r2.stdout)         # other statistics -> other features -> another calibrated threshold
    bad = subprocess.run([sys.executable, str(CLI), "--model_type", "robust", "--synthetic", "4"], capture_output=True, text=True,
                         timeout=600)
    assert bad.returncode != 0 and "RobustBench" in (bad.stderr + bad.stdout)
