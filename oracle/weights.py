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
