"""Drop-in mirror of the reference's attack.py (same names, arguments and error behaviour), with the
per-iteration work done by libadvb200 on the B200.

    reference                       here
    attack.py:10-17  normalize_batch    normalize_batch (torch; the native model fuses it into conv1)
    attack.py:20-34  fgsm_attack        fgsm_attack      -> advb200_input_grad + advb200_linf_step_f32
    attack.py:37-65  pgd_linf_attack    pgd_linf_attack  -> advb200_random_start_f32 + steps x (grad, step)
    attack.py:72-75  CWResult           CWResult
    attack.py:78-140 cw_l2_attack       cw_l2_attack     -> advb200_cw_init/step/finalize_f32

`model` is the torchvision ResNet-50 nn.Module the reference passes (packed once, cached) or a
NativeResNet50.  Inputs are never mutated; results are new detached tensors on x's device.
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass
from typing import Optional

import torch

from . import _lib
from .model import native_model


# The fused step (update inside the stem input-gradient epilogue, NativeResNet50.attack_step) is bit-identical
# but measured 3.5 % slower end to end (that epilogue is already the stem kernel's critical path), so the default
# is gradient + the HBM-rate update kernel; ADVB200_FUSED_STEP=1 opts in.
_FUSED_STEP = bool(os.environ.get("ADVB200_FUSED_STEP"))
# PGD's step loop runs natively (advb200_pgd_steps): one host call per step replaying a captured CUDA graph of the ~105
# kernel launches.  ADVB200_NO_NATIVE_LOOP=1 keeps the Python loop over input_grad + linf_step, ADVB200_NO_GRAPH=1 keeps
# the native loop but launches the kernels one by one.  All three give the same bits (tests/test_gpu_cw_kernels.py).
_NATIVE_LOOP = not os.environ.get("ADVB200_NO_NATIVE_LOOP")
_USE_GRAPH = not os.environ.get("ADVB200_NO_GRAPH")


def normalize_batch(x: torch.Tensor, mean: torch.Tensor, std: torch.Tensor) -> torch.Tensor:
    if mean.ndim != 1 or std.ndim != 1:
        raise ValueError("mean/std must be 1D tensors with shape [C]")
    if x.ndim != 4:
        raise ValueError("x must be a 4D tensor with shape [N, C, H, W]")
    mean = mean.to(device=x.device, dtype=x.dtype).view(1, -1, 1, 1)
    std = std.to(device=x.device, dtype=x.dtype).view(1, -1, 1, 1)
    return (x - mean) / std


def _check_common(x, mean, std):
    if mean.ndim != 1 or std.ndim != 1:
        raise ValueError("mean/std must be 1D tensors with shape [C]")
    if x.ndim != 4:
        raise ValueError("x must be a 4D tensor with shape [N, C, H, W]")


def _check_labels(y: torch.Tensor, n: int, classes: int = 1000) -> torch.Tensor:
    """int64 [N] labels on the host's word: F.cross_entropy / one_hot raise on a label outside [0, classes) in the
    reference (attack.py:31,114); here the check runs once per attack call (one tiny device read), not per step."""
    if y.ndim != 1 or y.shape[0] != n:
        raise ValueError("labels must have shape [N]")
    if n and bool(((y < 0) | (y >= classes)).any()):
        raise ValueError(f"labels must lie in [0, {classes})")
    return y


def _dev_vec(t: torch.Tensor, device) -> torch.Tensor:
    """mean / std moved to the device ONCE per attack call (the per-step calls then reuse the device tensor)."""
    return t.detach().to(device=device, dtype=torch.float32).contiguous() if t.ndim == 1 else t


def _linf_step(x_adv, x_orig, grad, alpha, eps, project):
    lib = _lib.load()
    rc = lib.advb200_linf_step_f32(_lib.ptr(x_adv), _lib.ptr(x_orig), _lib.ptr(grad), float(alpha), float(eps),
                                   x_adv.numel(), int(project), _lib.stream_ptr())
    _lib.check(rc, "advb200_linf_step_f32")


def fgsm_attack(model, x: torch.Tensor, y_true: torch.Tensor, *, eps: float, mean: torch.Tensor,
                std: torch.Tensor, inv_n_global: Optional[float] = None) -> torch.Tensor:
    _check_common(x, mean, std)
    nm = native_model(model)
    with torch.cuda.device(x.device):
        x_adv = x.detach().to(torch.float32).clone().contiguous()
        if x_adv.shape[0] == 0:
            return x_adv
        _check_labels(y_true, x_adv.shape[0])
        mean, std = _dev_vec(mean, x.device), _dev_vec(std, x.device)
        if nm.precision == "bf16" and _FUSED_STEP:
            nm.attack_step(x_adv, None, mean, std, y_true, alpha=eps, eps=eps, inv_n_global=inv_n_global)
        else:
            g = nm.input_grad(x_adv, mean, std, y_true, loss="ce", inv_n_global=inv_n_global)["grad"]
            _linf_step(x_adv, None, g, eps, eps, project=False)
    return x_adv


def pgd_linf_attack(model, x: torch.Tensor, y_true: torch.Tensor, *, eps: float, alpha: float, steps: int,
                    mean: torch.Tensor, std: torch.Tensor, random_start: bool = True,
                    inv_n_global: Optional[float] = None, noise: Optional[torch.Tensor] = None) -> torch.Tensor:
    """attack.py:37-65.  Extras beyond the reference signature (both optional): `inv_n_global` (gradient scale of a
    sharded batch) and `noise`, a caller-drawn U(-eps,eps) start tensor used instead of drawing one here — the parity
    harness injects the tensor the reference drew (CPU generator != CUDA Philox stream)."""
    _check_common(x, mean, std)
    if noise is not None and tuple(noise.shape) != tuple(x.shape):
        raise ValueError("noise must have the shape of x")
    nm = native_model(model)
    lib = _lib.load()
    with torch.cuda.device(x.device):
        x_orig = x.detach().to(torch.float32).contiguous()
        x_adv = torch.empty_like(x_orig)
        if x_orig.shape[0] == 0:
            return x_adv
        if random_start or noise is not None:
            if noise is None:
                noise = torch.empty_like(x_orig).uniform_(-eps, eps)      # attack.py:51, torch's generator
            noise = noise.detach().to(device=x_orig.device, dtype=torch.float32).contiguous()
            rc = lib.advb200_random_start_f32(_lib.ptr(x_adv), _lib.ptr(x_orig), _lib.ptr(noise),
                                              x_adv.numel(), _lib.stream_ptr())
            _lib.check(rc, "advb200_random_start_f32")
        else:
            x_adv.copy_(x_orig)
        n = x_orig.shape[0]
        _check_labels(y_true, n)
        mean, std = _dev_vec(mean, x.device), _dev_vec(std, x.device)
        y_dev = y_true.detach().to(device=x.device, dtype=torch.int64).contiguous()
        inv = 1.0 / n if inv_n_global is None else inv_n_global        # CE is a mean over the WHOLE batch (attack.py:58)
        # images are independent, so a batch larger than one workspace chunk runs all its steps chunk by chunk: the
        # engine keeps one launch plan per (batch size, workspace) and is not re-planned between steps
        grad = None
        for lo in range(0, n, nm.max_chunk):
            hi = min(n, lo + nm.max_chunk)
            xa, xo, yy = x_adv[lo:hi], x_orig[lo:hi], y_dev[lo:hi]
            if not (nm.precision == "bf16" and _FUSED_STEP) and _NATIVE_LOOP and tuple(xa.shape[1:]) == (3, 224, 224):
                # the whole loop of this chunk as one native call per step (captured CUDA graph, replayed)
                nm.pgd_steps(xa, xo, mean, std, yy, alpha=alpha, eps=eps, steps=int(steps), inv_n_global=inv, use_graph=_USE_GRAPH)
                continue
            if grad is None:
                grad = torch.empty((min(n, nm.max_chunk),) + tuple(x_orig.shape[1:]), dtype=torch.float32, device=x.device)
            g = grad[:hi - lo]
            for _ in range(int(steps)):
                if nm.precision == "bf16" and _FUSED_STEP:
                    # fused step: the sign step / projection / clamp run in the stem input-gradient kernel's epilogue
                    nm.attack_step(xa, xo, mean, std, yy, alpha=alpha, eps=eps, inv_n_global=inv)
                else:
                    nm.input_grad(xa, mean, std, yy, loss="ce", inv_n_global=inv, grad_out=g)
                    _linf_step(xa, xo, g, alpha, eps, project=True)
    return x_adv


@dataclass(frozen=True)
class CWResult:
    x_adv: torch.Tensor
    success: torch.Tensor


def cw_l2_attack(model, x: torch.Tensor, y_true: torch.Tensor, *, mean: torch.Tensor, std: torch.Tensor,
                 c: float = 1.0, kappa: float = 0.0, steps: int = 1000, lr: float = 1e-2,
                 targeted: bool = False, y_target: Optional[torch.Tensor] = None,
                 return_l2: bool = False):
    if targeted and y_target is None:
        raise ValueError("y_target must be provided when targeted=True")
    _check_common(x, mean, std)
    nm = native_model(model)
    lib = _lib.load()
    beta1, beta2, adam_eps = 0.9, 0.999, 1e-8          # torch.optim.Adam defaults (attack.py:101)
    with torch.cuda.device(x.device):
        xin = x.detach().to(torch.float32).contiguous()
        n = xin.shape[0]
        dev = xin.device
        if n == 0:
            return CWResult(x_adv=xin.clone(), success=torch.zeros((0,), dtype=torch.bool, device=dev))
        per = xin[0].numel()
        parts = int(lib.advb200_cw_parts(per))
        y_cmp = y_target if targeted else y_true
        _check_labels(y_cmp, n)
        y_cmp = y_cmp.detach().to(device=dev, dtype=torch.int64).contiguous()
        mean, std = _dev_vec(mean, dev), _dev_vec(std, dev)
        loss = "cw_targeted" if targeted else "cw"
        final = torch.empty_like(xin)
        best_l2 = torch.empty((n,), dtype=torch.float32, device=dev)
        best_success = torch.empty((n,), dtype=torch.uint8, device=dev)
        # CW's objective is a sum over images (attack.py:127), so chunks of the batch are independent problems: each
        # chunk runs all its steps on one launch plan and one set of state arrays
        chunk = min(n, nm.max_chunk)
        x0, w, m, v, x_adv, best_adv, grad = (torch.empty((chunk,) + tuple(xin.shape[1:]), dtype=torch.float32, device=dev)
                                              for _ in range(7))
        success = torch.empty((chunk,), dtype=torch.uint8, device=dev)
        l2_now = torch.empty((chunk,), dtype=torch.float32, device=dev)
        for lo in range(0, n, chunk):
            hi = min(n, lo + chunk)
            k = hi - lo
            # [k*parts floats] + k flag bytes (the step kernel keeps its `improved` flags in the tail)
            pa = torch.empty((k * parts + (k + 3) // 4 + 1,), dtype=torch.float32, device=dev)
            pb = torch.empty_like(pa)
            bl2, bs = best_l2[lo:hi], best_success[lo:hi]
            rc = lib.advb200_cw_init_f32(_lib.ptr(xin[lo:hi]), _lib.ptr(x0), _lib.ptr(w), _lib.ptr(m), _lib.ptr(v),
                                         _lib.ptr(x_adv), _lib.ptr(best_adv), _lib.ptr(bl2), _lib.ptr(bs), _lib.ptr(pa), k, per,
                                         _lib.stream_ptr())
            _lib.check(rc, "advb200_cw_init_f32")
            for t in range(1, int(steps) + 1):
                # forward + margin-loss input gradient at the current iterate; success flags of attack.py:121,124
                nm.input_grad(x_adv[:k], mean, std, y_cmp[lo:hi], loss=loss, c=c, kappa=kappa, grad_out=grad[:k],
                              success_out=success[:k])
                # Adam scalars in double on the host, like torch's single-tensor path
                step_size = lr / (1.0 - beta1 ** t)
                bc2_sqrt = math.sqrt(1.0 - beta2 ** t)
                rc = lib.advb200_cw_step_f32(_lib.ptr(w), _lib.ptr(m), _lib.ptr(v), _lib.ptr(x0), _lib.ptr(grad),
                                             _lib.ptr(x_adv), _lib.ptr(best_adv), _lib.ptr(bl2), _lib.ptr(bs),
                                             _lib.ptr(success), _lib.ptr(pa), _lib.ptr(pb), _lib.ptr(l2_now), k, per,
                                             step_size, bc2_sqrt, beta1, beta2, adam_eps, _lib.stream_ptr())
                _lib.check(rc, "advb200_cw_step_f32")
                pa, pb = pb, pa
            rc = lib.advb200_cw_finalize_f32(_lib.ptr(x_adv), _lib.ptr(best_adv), _lib.ptr(bs), _lib.ptr(final[lo:hi]), k,
                                             per, _lib.stream_ptr())
            _lib.check(rc, "advb200_cw_finalize_f32")
    res = CWResult(x_adv=final, success=best_success.to(torch.bool))
    if return_l2:
        return res, best_l2
    return res
