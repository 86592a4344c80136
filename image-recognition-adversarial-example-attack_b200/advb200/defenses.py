"""Drop-in mirror of the defense / evaluation functions of the reference's defense_experiments.py.

    reference (defense_experiments.py)                 here
    :177-179 _defense_smoothing                         advb200_smooth3x3_f32   (bit-exact)
    :182-185 _defense_quantization                      advb200_quantize_f32    (bit-exact)
    :188-207 _jpeg_compress_batch                       advb200_jpeg_roundtrip_f32 (libjpeg sample pipeline on the device)
    :210-217 defend_input                               advb200_defend_input_f32 (one fused pass) [+ the JPEG stage if switched on]
    :221-257 _layer3_features                           advb200_forward(stop_at_layer3, score_out)
    :273-285 is_adversarial_by_feature                  score > threshold
    :288-345 calibrate_feature_threshold                quantile tail on native scores
    :534-576 run_attack                                 dispatch to the native attacks
    :579-646 evaluate_defenses                          same dict (batch 1) + a batched variant that
                                                        returns the int64[7] counter vector
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib
from .attack import cw_l2_attack, fgsm_attack, pgd_linf_attack
from .model import native_model

# module-level JPEG switches, as in the reference (defense_experiments.py:173-174; set by --use_jpeg / --jpeg_quality)
_DEFENSE_USE_JPEG = False
_DEFENSE_JPEG_QUALITY = 75

COUNTER_KEYS = ("clean_correct", "attack_success", "defense_preproc_success", "detector_flags_clean",
                "detector_flags_adv", "detector_attack_success", "count")


def _as_planes(x: torch.Tensor):
    if x.ndim != 4:
        raise ValueError("x must be a 4D tensor with shape [N, C, H, W]")
    if not x.is_cuda:
        raise ValueError("x must be a CUDA tensor (no CPU fallback)")
    xc = x.detach().to(torch.float32).contiguous()
    n, c, h, w = xc.shape
    return xc, n * c, h, w


def _defense_smoothing(x: torch.Tensor) -> torch.Tensor:
    xc, planes, h, w = _as_planes(x)
    out = torch.empty_like(xc)
    with torch.cuda.device(xc.device):
        rc = _lib.load().advb200_smooth3x3_f32(_lib.ptr(xc), _lib.ptr(out), planes, h, w, _lib.stream_ptr())
    _lib.check(rc, "advb200_smooth3x3_f32")
    return out


def _defense_median_smoothing(x: torch.Tensor) -> torch.Tensor:
    """3x3 median smoothing (stride 1, zero pad 1) — not in the reference, whose `_defense_smoothing` is the 3x3
    mean (defense_experiments.py:177-179); offered as an extra defense with the same calling convention."""
    xc, planes, h, w = _as_planes(x)
    out = torch.empty_like(xc)
    with torch.cuda.device(xc.device):
        rc = _lib.load().advb200_median3x3_f32(_lib.ptr(xc), _lib.ptr(out), planes, h, w, _lib.stream_ptr())
    _lib.check(rc, "advb200_median3x3_f32")
    return out


def _defense_quantization(x: torch.Tensor, levels: int = 16) -> torch.Tensor:
    if not x.is_cuda:
        raise ValueError("x must be a CUDA tensor (no CPU fallback)")
    xc = x.detach().to(torch.float32).contiguous()
    out = torch.empty_like(xc)
    with torch.cuda.device(xc.device):
        rc = _lib.load().advb200_quantize_f32(_lib.ptr(xc), _lib.ptr(out), xc.numel(), int(levels),
                                              _lib.stream_ptr())
    _lib.check(rc, "advb200_quantize_f32")
    return out


def _jpeg_compress_batch(x: torch.Tensor, quality: int) -> torch.Tensor:
    """defense_experiments.py:188-207 without the host round trip: the PIL/libjpeg sample pipeline (4:2:0, islow
    DCT, quality-scaled tables, fancy upsampling) runs on the device, bit-exact with the reference's PIL path."""
    if x.ndim != 4 or x.shape[1] != 3:
        raise ValueError("x must be a 4D RGB tensor with shape [N, 3, H, W]")
    if not x.is_cuda:
        raise ValueError("x must be a CUDA tensor (no CPU fallback)")
    xc = x.detach().to(torch.float32).contiguous()
    n, _, h, w = xc.shape
    out = torch.empty_like(xc)
    if n == 0 or h == 0 or w == 0:
        return out.to(x.dtype)
    lib = _lib.load()
    with torch.cuda.device(xc.device):
        nbytes = int(lib.advb200_jpeg_scratch_bytes(n, h, w))
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=xc.device)
        rc = lib.advb200_jpeg_roundtrip_f32(_lib.ptr(xc), _lib.ptr(out), n, h, w, int(quality), _lib.ptr(scratch),
                                            nbytes, _lib.stream_ptr())
    _lib.check(rc, "advb200_jpeg_roundtrip_f32")
    return out.to(x.dtype)


def defend_input(x: torch.Tensor) -> torch.Tensor:
    xc, planes, h, w = _as_planes(x)
    out = torch.empty_like(xc)
    with torch.cuda.device(xc.device):
        rc = _lib.load().advb200_defend_input_f32(_lib.ptr(xc), _lib.ptr(out), planes, h, w, 16,
                                                  _lib.stream_ptr())
    _lib.check(rc, "advb200_defend_input_f32")
    if _DEFENSE_USE_JPEG:                      # defense_experiments.py:215-216 (the trailing clamp is a no-op)
        out = _jpeg_compress_batch(out, _DEFENSE_JPEG_QUALITY)
    return out


def _layer3_features(model, x: torch.Tensor, mean: torch.Tensor, std: torch.Tensor) -> torch.Tensor:
    if mean.ndim != 1 or std.ndim != 1:
        raise ValueError("mean/std must be 1D tensors with shape [C]")
    nm = native_model(model)
    return nm.forward(x, mean, std, want_logits=False, want_score=True, stop_at_layer3=True)["score"]


def is_adversarial_by_feature(model, x, mean, std, threshold: float) -> torch.Tensor:
    return _layer3_features(model, x, mean, std) > float(threshold)


def calibrate_threshold_from_scores(scores: torch.Tensor, quantile: float = 0.95) -> float:
    """Scalar tail of calibrate_feature_threshold (defense_experiments.py:327-345)."""
    if scores.numel() == 0:
        return 5.0
    thr = float(torch.quantile(scores.detach().float().cpu(), float(quantile)).item())
    if thr > 50:
        return thr * 0.5
    return max(thr, 1.0)


def calibrate_feature_threshold(model, images, *, mean, std, n: int = 100, quantile: float = 0.95,
                                device=None, transform=None) -> float:
    """`images` is a [M,3,224,224] tensor (synthetic batches) or a list of image paths (then
    `transform` and `device` are used exactly as in the reference)."""
    if isinstance(images, torch.Tensor):
        num = min(int(n), images.shape[0])
        if num <= 0:
            raise ValueError("no calibration images available")
        scores = _layer3_features(model, images[:num], mean, std)
        return calibrate_threshold_from_scores(scores, quantile)
    num = min(int(n), len(images))
    if num <= 0:
        raise ValueError("no calibration images available")
    from PIL import Image
    xs = []
    for p in images[:num]:
        try:
            xs.append(transform(Image.open(p).convert("RGB")).unsqueeze(0))
        except Exception as e:  # same tolerance as the reference loop (:310-320)
            print(f"calibration image failed {getattr(p, 'name', p)}: {e}")
    if not xs:
        return 5.0
    x = torch.cat(xs, 0).to(device)
    return calibrate_threshold_from_scores(_layer3_features(model, x, mean, std), quantile)


def run_attack(attack_name, model, x, y_true, mean, std, eps, alpha, steps, cw_c, cw_kappa, cw_steps, cw_lr,
               inv_n_global: Optional[float] = None, pgd_noise: Optional[torch.Tensor] = None):
    if attack_name == "fgsm":
        return fgsm_attack(model, x, y_true, eps=eps, mean=mean, std=std, inv_n_global=inv_n_global)
    if attack_name == "pgd":
        return pgd_linf_attack(model, x, y_true, eps=eps, alpha=alpha, steps=steps, mean=mean, std=std,
                               inv_n_global=inv_n_global, noise=pgd_noise)
    return cw_l2_attack(model, x, y_true, mean=mean, std=std, c=cw_c, kappa=cw_kappa, steps=cw_steps, lr=cw_lr,
                        targeted=False, y_target=None).x_adv


def evaluate_defenses_batched(model, x, y_true, mean, std, attack_name, eps, alpha, steps, cw_c, cw_kappa,
                              cw_steps, cw_lr, detector_threshold, *, inv_n_global: Optional[float] = None,
                              return_flags: bool = False, x_adv: Optional[torch.Tensor] = None,
                              pgd_noise: Optional[torch.Tensor] = None, return_details: bool = False):
    """Vectorised evaluate_defenses: per-image outcomes equal the reference's batch-1 loop
    (every image is independent; CE's batch mean only rescales the gradient).  Returns the int64[7]
    counter vector {COUNTER_KEYS} on x's device (and the [N,6] per-image flags if asked).
    With inv_n_global=None the gradient scale is 1 (what each image sees in a batch-1 loop)."""
    nm = native_model(model)
    lib = _lib.load()
    n = x.shape[0]
    dev = x.device
    scale = 1.0 if inv_n_global is None else inv_n_global
    clean = nm.forward(x, mean, std, want_logits=False, want_preds=True, want_score=True)
    if x_adv is None:
        x_adv = run_attack(attack_name, nm, x, y_true, mean, std, eps, alpha, steps, cw_c, cw_kappa, cw_steps,
                           cw_lr, inv_n_global=scale, pgd_noise=pgd_noise)
    adv = nm.forward(x_adv, mean, std, want_logits=False, want_preds=True, want_score=True)
    x_def = defend_input(x_adv)
    dfd = nm.forward(x_def, mean, std, want_logits=False, want_preds=True)
    counters = torch.zeros((7,), dtype=torch.int64, device=dev)
    flags = torch.empty((n, 6), dtype=torch.uint8, device=dev) if (return_flags or return_details) else None
    y = y_true.detach().to(device=dev, dtype=torch.int64).contiguous()
    with torch.cuda.device(dev):
        rc = lib.advb200_eval_counters(_lib.ptr(clean["preds"]), _lib.ptr(adv["preds"]), _lib.ptr(dfd["preds"]),
                                       _lib.ptr(clean["score"]), _lib.ptr(adv["score"]),
                                       float(detector_threshold), _lib.ptr(y), n, _lib.ptr(counters),
                                       _lib.ptr(flags), _lib.stream_ptr())
    _lib.check(rc, "advb200_eval_counters")
    if return_details:       # per-image evidence for the parity reports: flags, detector scores, the three predictions
        return counters, flags, {"score_clean": clean["score"], "score_adv": adv["score"], "pred_clean": clean["preds"],
                                 "pred_adv": adv["preds"], "pred_def": dfd["preds"]}
    if return_flags:
        return counters, flags
    return counters


def evaluate_defenses(model, x, y_true, mean, std, attack_name, eps, alpha, steps, cw_c, cw_kappa, cw_steps,
                      cw_lr, detector_threshold) -> dict:
    """Reference signature and return dict (defense_experiments.py:579-646); batch 1 like the reference."""
    if x.shape[0] != 1:
        raise ValueError("evaluate_defenses mirrors the reference's batch-1 contract; "
                         "use evaluate_defenses_batched for N > 1")
    counters = evaluate_defenses_batched(model, x, y_true, mean, std, attack_name, eps, alpha, steps, cw_c,
                                         cw_kappa, cw_steps, cw_lr, detector_threshold).cpu().tolist()
    out = {k: int(v) for k, v in zip(COUNTER_KEYS, counters) if k != "count"}
    out["clean_correct"] = bool(out["clean_correct"])
    return out
