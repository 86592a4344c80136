"""Generates tests/golden/* by importing and running the REFERENCE ITSELF from /root/reference
(read-only).  Runs only in the build container (the GPU box has no /root/reference); the outputs
are committed so the tests can check both the oracle restatement and the CUDA path against them.

    python oracle/gen_golden.py            # ~2 min on 8 cores
"""
from __future__ import annotations

import contextlib
import copy
import io
import json
import runpy
import sys
import types
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
REF = Path("/root/reference")
OUT = ROOT / "tests" / "golden"
sys.path.insert(0, str(ROOT))
sys.dont_write_bytecode = True

from oracle import weights  # noqa: E402


def import_reference():
    sys.path.insert(0, str(REF))
    import attack as ref_attack  # noqa
    # defense_experiments.py imports matplotlib at module level (line 19); stub it out
    for name in ("matplotlib", "matplotlib.pyplot"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    import defense_experiments as ref_def  # noqa
    return ref_attack, ref_def


def tiny_model(wt):
    return lambda z: z.flatten(1) @ wt


def main():
    OUT.mkdir(parents=True, exist_ok=True)
    ref_attack, ref_def = import_reference()
    meta = {"torch": torch.__version__, "threads": torch.get_num_threads()}
    mean, std = weights.mean_std()

    # ---- 1. attack loops on a tiny linear classifier (exercises every line of attack.py) ----
    g = torch.Generator().manual_seed(11)
    x = torch.rand(3, 3, 4, 4, generator=g)
    wt = torch.randn(48, 5, generator=g)
    y = torch.tensor([0, 3, 1])
    yt = torch.tensor([2, 2, 4])
    eps, alpha = 8 / 255, 2 / 255
    tm = tiny_model(wt)
    fg = ref_attack.fgsm_attack(tm, x, y, eps=eps, mean=mean, std=std)
    pg0 = ref_attack.pgd_linf_attack(tm, x, y, eps=eps, alpha=alpha, steps=7, mean=mean, std=std, random_start=False)
    torch.manual_seed(5)
    pg1 = ref_attack.pgd_linf_attack(tm, x, y, eps=eps, alpha=alpha, steps=7, mean=mean, std=std)
    cwu = ref_attack.cw_l2_attack(tm, x, y, mean=mean, std=std, c=2.0, kappa=0.1, steps=25, lr=0.05)
    cwt = ref_attack.cw_l2_attack(tm, x, y, mean=mean, std=std, c=2.0, kappa=0.0, steps=25, lr=0.05, targeted=True,
                                  y_target=yt)
    np.savez(OUT / "tiny_linear.npz", x=x.numpy(), wt=wt.numpy(), y=y.numpy(), yt=yt.numpy(), eps=eps, alpha=alpha,
             fgsm=fg.numpy(), pgd_nostart=pg0.numpy(), pgd_seed5=pg1.numpy(), cwu_x=cwu.x_adv.numpy(),
             cwu_s=cwu.success.numpy(), cwt_x=cwt.x_adv.numpy(), cwt_s=cwt.success.numpy())

    # ---- 2. defenses on a small tensor ---------------------------------------------------------
    g = torch.Generator().manual_seed(12)
    d = torch.rand(2, 3, 9, 13, generator=g) * 1.3 - 0.15
    d[0, 0, 0, :6] = torch.tensor([0.5 / 15, 1.5 / 15, 2.5 / 15, 0.0, 1.0, 7.5 / 15])   # rounding ties
    np.savez(OUT / "defense_small.npz", x=d.numpy(), smooth=ref_def._defense_smoothing(d).numpy(),
             quant16=ref_def._defense_quantization(d, 16).numpy(), quant4=ref_def._defense_quantization(d, 4).numpy(),
             defend=ref_def.defend_input(d).numpy())

    # ---- 3. ResNet-50 logits / input gradients, fp32 and fp64 ----------------------------------
    model = weights.seeded_resnet50(0)
    xs = weights.synthetic_images(4)
    x2 = xs[:2]
    xi = x2.clone().requires_grad_(True)
    logits = model(ref_attack.normalize_batch(xi, mean, std))
    y2 = logits.argmax(1)
    (grad,) = torch.autograd.grad(torch.nn.functional.cross_entropy(logits, y2), xi)
    m64 = copy.deepcopy(model).double()
    xi64 = x2.double().clone().requires_grad_(True)
    l64 = m64(ref_attack.normalize_batch(xi64, mean.double(), std.double()))
    (g64,) = torch.autograd.grad(torch.nn.functional.cross_entropy(l64, y2), xi64)
    np.savez_compressed(OUT / "resnet_n2.npz", logits=logits.detach().numpy(), logits64=l64.detach().numpy(),
                        y=y2.numpy(), grad=grad.numpy(), grad64=g64.float().numpy())
    meta["ref_fp32_vs_fp64"] = {
        "logits_rel": float((logits.double() - l64).norm() / l64.norm()),
        "grad_rel": float((grad.double() - g64).norm() / g64.norm()),
    }

    # ---- 4. attacks on the ResNet (N=2) -----------------------------------------------------
    fg = ref_attack.fgsm_attack(model, x2, y2, eps=eps, mean=mean, std=std)
    torch.manual_seed(7)
    pg = ref_attack.pgd_linf_attack(model, x2, y2, eps=eps, alpha=alpha, steps=10, mean=mean, std=std)
    with torch.no_grad():
        lf = model(ref_attack.normalize_batch(fg, mean, std))
        lp = model(ref_attack.normalize_batch(pg, mean, std))
    cwu = ref_attack.cw_l2_attack(model, x2, y2, mean=mean, std=std, c=1.0, kappa=0.0, steps=20, lr=0.01)
    cwt = ref_attack.cw_l2_attack(model, x2, y2, mean=mean, std=std, c=1.0, kappa=0.0, steps=20, lr=0.01,
                                  targeted=True, y_target=torch.full((2,), 805))
    with torch.no_grad():
        lcu = model(ref_attack.normalize_batch(cwu.x_adv, mean, std))
        lct = model(ref_attack.normalize_batch(cwt.x_adv, mean, std))
    l2 = lambda a: (a - x2.clamp(0, 1)).flatten(1).pow(2).sum(1).numpy()
    np.savez_compressed(OUT / "attacks_n2.npz", fgsm_logits=lf.numpy(), pgd_logits=lp.numpy(),
                        pgd_linf=(pg - x2).abs().flatten(1).amax(1).numpy(), pgd_x=pg.numpy().astype(np.float32),
                        cwu_success=cwu.success.numpy(), cwu_l2=l2(cwu.x_adv), cwu_logits=lcu.numpy(),
                        cwt_success=cwt.success.numpy(), cwt_l2=l2(cwt.x_adv), cwt_logits=lct.numpy())

    # ---- 5. detector scores, threshold, evaluate_defenses (batch 1, as the reference runs it) ---
    scores = torch.cat([ref_def._layer3_features(model, xs[i:i + 1], mean, std) for i in range(4)])
    q = float(torch.quantile(scores, 0.95).item())
    thr = q * 0.5 if q > 50 else max(q, 1.0)          # defense_experiments.py:338-345
    evals = {}
    for attack_name in ("fgsm", "pgd"):
        for e in (4 / 255, 8 / 255):
            rows = []
            for i in range(2):
                xi1 = xs[i:i + 1]
                yi = model(ref_attack.normalize_batch(xi1, mean, std)).argmax(1)
                torch.manual_seed(100 + i)
                r = ref_def.evaluate_defenses(model, xi1, yi, mean, std, attack_name, float(e), alpha, 10, 1.0, 0.0,
                                              20, 0.01, thr)
                rows.append({k: int(v) for k, v in r.items()})
            evals[f"{attack_name}_{round(e * 255)}"] = rows
    meta["detector"] = {"scores": scores.tolist(), "threshold": thr, "evaluate_defenses": evals}

    # ---- 6. config 1: the reference CLI on picture/example.jpg with seeded random-init weights ---
    import torchvision
    real_resnet50 = torchvision.models.resnet50

    def patched(*a, **k):
        torch.manual_seed(0)
        return real_resnet50(weights=None)

    torchvision.models.resnet50 = patched
    argv = sys.argv
    sys.argv = ["ResNet.py", str(REF / "picture" / "example.jpg"), "--attack", "fgsm", "--eps", "0.031372549",
                "--topk", "5"]
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        ns = runpy.run_path(str(REF / "ResNet.py"), run_name="__main__")
    sys.argv = argv
    torchvision.models.resnet50 = real_resnet50
    meta["cli_config1_stdout"] = buf.getvalue().splitlines()
    from PIL import Image
    xcli = ns["transform"](Image.open(REF / "picture" / "example.jpg").convert("RGB"))
    u8 = (xcli * 255).round().to(torch.uint8)
    assert torch.equal(u8.float() / 255, xcli)
    np.save(OUT / "example_224_u8.npy", u8.numpy())

    (OUT / "meta.json").write_text(json.dumps(meta, indent=1))
    print(json.dumps(meta, indent=1)[:3000])


if __name__ == "__main__":
    main()
