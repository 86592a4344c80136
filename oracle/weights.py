"""Shared seeded inputs of every parity test: fixed random-init ResNet-50 weights
(pretrained checkpoints are unavailable offline), synthetic U[0,1) images, ImageNet mean/std."""
from __future__ import annotations

import torch

MEAN = (0.485, 0.456, 0.406)   # ResNet.py:23, defense_experiments.py:666
STD = (0.229, 0.224, 0.225)    # ResNet.py:24, defense_experiments.py:667


def seeded_resnet50(seed: int = 0) -> torch.nn.Module:
    """torchvision.models.resnet50(weights=None).eval() after torch.manual_seed(seed) — the model
    ResNet.py:13 / defense_experiments.py:690 would build, minus the download."""
    import torchvision
    torch.manual_seed(seed)
    return torchvision.models.resnet50(weights=None).eval()


def synthetic_images(n: int, seed: int = 1234) -> torch.Tensor:
    return torch.rand(n, 3, 224, 224, generator=torch.Generator().manual_seed(seed))


def mean_std():
    return torch.tensor(MEAN), torch.tensor(STD)


def pgd_noise(shape, eps: float, seed: int) -> torch.Tensor:
    """The tensor attack.py:51 draws when called right after torch.manual_seed(seed)."""
    torch.manual_seed(seed)
    return torch.empty(shape).uniform_(-eps, eps)


# ---------------------------------------------------------------------------------------------------------------------
# A second, non-degenerate fixture model.  The seeded random-init model above has identity BatchNorm statistics
# (scale 1/sqrt(1+eps), shift 0 everywhere), activations that grow to |a| ~ 300 and the same predicted class for every
# image.  `conditioned_resnet50` is the same architecture and the same seeded convolution / fc weights
# with every BatchNorm given non-trivial parameters, so the folded scale AND shift of every conv are exercised, and with
# running statistics taken from one calibration pass so activations stay O(1) at every depth.
# ---------------------------------------------------------------------------------------------------------------------
_GOLDEN = __import__("pathlib").Path(__file__).resolve().parent.parent / "tests" / "golden"


def structured_images(n: int, seed: int = 77) -> torch.Tensor:
    """Synthetic images with spatial structure (a low-frequency colour field at three scales, an oriented sinusoid, fine
    noise), quantised to k/255 like decoded photographs — quantisation also makes the tensor identical on every host,
    whatever the last-bit behaviour of the interpolation kernels."""
    import math
    import torch.nn.functional as F
    g = torch.Generator().manual_seed(seed)
    out = torch.zeros(n, 3, 224, 224)
    for s, amp in ((4, 0.30), (14, 0.18), (56, 0.10)):
        coarse = 0.5 * torch.randn(n, 3, s, s, generator=g) + 0.8 * torch.randn(n, 1, s, s, generator=g)
        out += amp * F.interpolate(coarse, size=(224, 224), mode="bilinear", align_corners=False)
    yy, xx = torch.meshgrid(torch.arange(224.0), torch.arange(224.0), indexing="ij")
    th = (torch.rand(n, generator=g) * math.pi).view(n, 1, 1)
    fr = (0.05 + 0.25 * torch.rand(n, generator=g)).view(n, 1, 1)
    ph = (torch.rand(n, generator=g) * 2 * math.pi).view(n, 1, 1)
    wave = torch.sin(fr * (xx * torch.cos(th) + yy * torch.sin(th)) + ph)
    out += 0.12 * wave.unsqueeze(1) * torch.randn(n, 3, 1, 1, generator=g)
    out += 0.03 * torch.randn(n, 3, 224, 224, generator=g)
    out += torch.rand(n, 3, 1, 1, generator=g) * 0.5 + 0.25
    return torch.round(out.clamp(0, 1) * 255) / 255


def mixed_images(n: int, seed: int = 77) -> torch.Tensor:
    """Even slots structured, odd slots U[0,1) noise — the image set of the conditioned-model fixtures."""
    s = structured_images((n + 1) // 2, seed)
    u = synthetic_images(n // 2, seed + 1) if n > 1 else s[:0]
    out = torch.empty(n, 3, 224, 224)
    out[0::2] = s
    out[1::2] = u
    return out


def _bn_modules(model):
    return [(name, mod) for name, mod in model.named_modules() if isinstance(mod, torch.nn.BatchNorm2d)]


def conditioned_resnet50(seed: int = 0, stats_file=None) -> torch.nn.Module:
    """seeded_resnet50(seed) with non-identity BatchNorm everywhere: gamma ~ U[0.5,1.5] (U[0.05,0.25] for bn3, the last
    BatchNorm of each residual branch — a random-init BN network's input gradient otherwise explodes with depth, which
    is exactly the ill-conditioning the fixture is meant to remove), beta ~ U[-0.2,0.2], running_mean / running_var from
    one training-mode calibration pass over 16 structured + 16 noise images.  The calibrated statistics are committed
    (tests/golden/conditioned_bn_seed<seed>.npz) so every host builds bit-identical weights; if the file is absent
    they are computed."""
    import numpy as np
    m = seeded_resnet50(seed)
    gen = torch.Generator().manual_seed(1000 + seed)
    bns = _bn_modules(m)
    for name, bn in bns:
        lo, hi = (0.05, 0.25) if name.endswith("bn3") else (0.5, 1.5)
        bn.weight.data = torch.rand(bn.num_features, generator=gen) * (hi - lo) + lo
        bn.bias.data = torch.rand(bn.num_features, generator=gen) * 0.4 - 0.2
    path = _GOLDEN / f"conditioned_bn_seed{seed}.npz" if stats_file is None else stats_file
    if path is not None and __import__("pathlib").Path(path).is_file():
        st = np.load(path)
        for name, bn in bns:
            bn.running_mean.data = torch.from_numpy(st[name + ".mean"]).clone()
            bn.running_var.data = torch.from_numpy(st[name + ".var"]).clone()
        return m.eval()
    for _, bn in bns:
        bn.momentum = 1.0            # running statistics := the calibration batch's statistics
    calib = mixed_images(32, seed=4321)
    mean, std = mean_std()
    m.train()
    with torch.no_grad():
        m((calib - mean.view(1, 3, 1, 1)) / std.view(1, 3, 1, 1))
    for _, bn in bns:
        bn.momentum = 0.1
    return m.eval()


def save_conditioned_stats(model, seed: int = 0):
    import numpy as np
    np.savez(_GOLDEN / f"conditioned_bn_seed{seed}.npz",
             **{f"{name}.{k}": getattr(bn, "running_" + k).numpy() for name, bn in _bn_modules(model) for k in ("mean", "var")})


MODELS = {"seed0": seeded_resnet50, "cond": conditioned_resnet50}
