"""Restatement of the reference's attack loops (attack.py) with explicit arithmetic.

Each function cites the reference lines it follows.  `model` is any callable mapping a normalised
[N,3,H,W] tensor to logits.  Differences from a verbatim copy are deliberate and visible: Adam is
written out (torch/optim/adam.py single-tensor path), the PGD start noise can be
injected, and the CW loop reports the per-image L2 it tracked.
"""
from __future__ import annotations

import math
from typing import Callable, Optional, Tuple

import torch
import torch.nn.functional as F


def normalize(x, mean, std):
    """attack.py:10-17."""
    if mean.ndim != 1 or std.ndim != 1:
        raise ValueError("mean/std must be 1D tensors with shape [C]")
    if x.ndim != 4:
        raise ValueError("x must be a 4D tensor with shape [N, C, H, W]")
    return (x - mean.to(x).view(1, -1, 1, 1)) / std.to(x).view(1, -1, 1, 1)


def ce_input_grad(model: Callable, x, y, mean, std) -> Tuple[torch.Tensor, torch.Tensor]:
    """attack.py:29-32 / 56-59: mean-reduced CE, gradient w.r.t. the un-normalised input only."""
    xi = x.detach().clone().requires_grad_(True)
    logits = model(normalize(xi, mean, std))
    loss = F.cross_entropy(logits, y)
    (g,) = torch.autograd.grad(loss, xi, only_inputs=True)
    return g, logits.detach()


def linf_update(x_adv, x_orig, grad, alpha, eps, project=True):
    """attack.py:61-63 (project) / attack.py:33-34 (FGSM: no projection)."""
    s = (grad > 0).to(x_adv.dtype) - (grad < 0).to(x_adv.dtype)
    out = x_adv + alpha * s
    if project:
        out = torch.max(torch.min(out, x_orig + eps), x_orig - eps)
    return out.clamp(0.0, 1.0)


def fgsm(model, x, y, *, eps, mean, std):
    """attack.py:20-34."""
    g, _ = ce_input_grad(model, x, y, mean, std)
    return linf_update(x.detach(), None, g, eps, eps, project=False)


def pgd_linf(model, x, y, *, eps, alpha, steps, mean, std, noise: Optional[torch.Tensor] = None):
    """attack.py:37-65; `noise` is the U(-eps,eps) start tensor (None = random_start=False)."""
    x_orig = x.detach()
    x_adv = (x_orig + noise).clamp(0.0, 1.0) if noise is not None else x_orig.clone()
    for _ in range(int(steps)):
        g, _ = ce_input_grad(model, x_adv, y, mean, std)
        x_adv = linf_update(x_adv, x_orig, g, alpha, eps, project=True)
    return x_adv


def adam_update(w, m, v, g, t, lr, b1=0.9, b2=0.999, adam_eps=1e-8):
    """torch.optim.Adam single-tensor update at step t >= 1 (betas .9/.999, eps 1e-8, no weight decay / amsgrad;
    torch/optim/adam.py): the bias corrections are Python doubles applied as fp32 scalars."""
    m = m + (g - m) * (1.0 - b1)
    v = v * b2 + (1.0 - b2) * g * g
    step_size = lr / (1.0 - b1 ** t)
    denom = v.sqrt() / math.sqrt(1.0 - b2 ** t) + adam_eps
    return w - step_size * (m / denom), m, v


def cw_l2(model, x, y_true, *, mean, std, c=1.0, kappa=0.0, steps=1000, lr=1e-2, targeted=False, y_target=None):
    """attack.py:78-140 with torch.optim.Adam written out.  Returns (x_adv, success, best_l2)."""
    if targeted and y_target is None:
        raise ValueError("y_target must be provided when targeted=True")
    x0 = x.detach().clamp(0.0, 1.0)
    eps = 1e-6
    t0 = (x0 * (1.0 - 2.0 * eps) + eps) * 2.0 - 1.0
    w = 0.5 * (torch.log1p(t0) - torch.log1p(-t0))                       # attack.py:68-69, 95-98
    m = torch.zeros_like(w)
    v = torch.zeros_like(w)
    b1, b2, adam_eps = 0.9, 0.999, 1e-8
    n = x0.shape[0]
    best_adv = x0.clone()
    best_l2 = torch.full((n,), float("inf"), dtype=x0.dtype)
    best_success = torch.zeros((n,), dtype=torch.bool)
    y_cmp = y_target if targeted else y_true
    for t in range(1, int(steps) + 1):
        wv = w.detach().clone().requires_grad_(True)
        x_adv = 0.5 * (torch.tanh(wv) + 1.0)                              # :110
        logits = model(normalize(x_adv, mean, std))
        onehot = F.one_hot(y_cmp, num_classes=logits.shape[1]).to(logits.dtype)
        real = (logits * onehot).sum(dim=1)                               # :116
        other = (logits - 1e4 * onehot).amax(dim=1)                       # :117
        if targeted:
            f = torch.clamp(other - real + kappa, min=0.0)
            success = logits.argmax(dim=1) == y_cmp
        else:
            f = torch.clamp(real - other + kappa, min=0.0)
            success = logits.argmax(dim=1) != y_cmp
        l2 = (x_adv - x0).view(n, -1).pow(2).sum(dim=1)                   # :126
        loss = (l2 + c * f).sum()                                         # :127
        (g,) = torch.autograd.grad(loss, wv)
        w, m, v = adam_update(w, m, v, g, t, lr, b1, b2, adam_eps)
        improved = success & (l2.detach() < best_l2)                      # :133-137
        best_l2 = torch.where(improved, l2.detach(), best_l2)
        best_success = best_success | improved
        best_adv = torch.where(improved.view(-1, 1, 1, 1), x_adv.detach(), best_adv)
    final = torch.where(best_success.view(-1, 1, 1, 1), best_adv, 0.5 * (torch.tanh(w) + 1.0))   # :139
    return final, best_success, best_l2
