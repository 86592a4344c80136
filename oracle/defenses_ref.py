"""Restatement of the reference's defenses and evaluation (defense_experiments.py) — numpy for
the bit-exact elementwise parts, torch for the detector statistics."""
from __future__ import annotations

import numpy as np
import torch

from . import attacks_ref


def smooth3x3(x: np.ndarray) -> np.ndarray:
    """defense_experiments.py:177-179 = F.avg_pool2d(x, 3, 1, 1): fp32 row-major sum of the nine
    taps (zero padding contributes +0), then TRUE division by 9."""
    x = np.asarray(x, dtype=np.float32)
    n, c, h, w = x.shape
    p = np.zeros((n, c, h + 2, w + 2), dtype=np.float32)
    p[:, :, 1:-1, 1:-1] = x
    s = np.zeros_like(x)
    for dr in range(3):
        for dc in range(3):
            s = (s + p[:, :, dr:dr + h, dc:dc + w]).astype(np.float32)
    return (s / np.float32(9.0)).astype(np.float32)


def median3x3(x: np.ndarray) -> np.ndarray:
    """3x3 median, stride 1, zero pad 1 (the window and padding of defense_experiments.py:177-179 with the median in
    place of the mean).  Not a reference function: the oracle for the extra median-smoothing defense; pinned against
    scipy.ndimage.median_filter(mode="constant") in tests/test_oracle_golden.py."""
    x = np.asarray(x, dtype=np.float32)
    n, c, h, w = x.shape
    p = np.zeros((n, c, h + 2, w + 2), dtype=np.float32)
    p[:, :, 1:-1, 1:-1] = x
    taps = np.stack([p[:, :, dr:dr + h, dc:dc + w] for dr in range(3) for dc in range(3)])
    return np.sort(taps, axis=0)[4]


def quantize(x: np.ndarray, levels: int = 16) -> np.ndarray:
    """defense_experiments.py:182-185: round-half-even of clamp(x,0,1)*(levels-1), TRUE division."""
    x = np.asarray(x, dtype=np.float32)
    lm1 = np.float32(levels - 1)
    return (np.rint((np.clip(x, 0.0, 1.0) * lm1).astype(np.float32)) / lm1).astype(np.float32)


def defend_input(x: np.ndarray, use_jpeg: bool = False, jpeg_quality: int = 75) -> np.ndarray:
    """defense_experiments.py:210-217; the JPEG branch (:215-216) is off by default (:173)."""
    y = quantize(smooth3x3(np.clip(np.asarray(x, dtype=np.float32), 0.0, 1.0)), 16)
    if use_jpeg:
        from . import jpeg_ref
        y = jpeg_ref.jpeg_compress_batch(y, jpeg_quality)
    return np.clip(y, 0.0, 1.0)


def layer3_score(feat: torch.Tensor) -> torch.Tensor:
    """defense_experiments.py:245-255 on the layer3 feature map [N,1024,14,14]."""
    l2 = feat.pow(2).sum(dim=[1, 2, 3]).sqrt() / feat.shape[1]
    var = feat.var(dim=[2, 3]).mean(dim=1)
    return torch.clamp(l2 + var * 0.1, 0, 100)


def calibrate_threshold(scores: torch.Tensor, quantile: float = 0.95) -> float:
    """Scalar tail of calibrate_feature_threshold (defense_experiments.py:327-345)."""
    if scores.numel() == 0:
        return 5.0
    thr = float(torch.quantile(scores, float(quantile)).item())
    return thr * 0.5 if thr > 50 else max(thr, 1.0)


def evaluate_defenses(model, x, y_true, mean, std, attack_name, eps, alpha, steps, threshold, *, noise=None,
                      cw=None):
    """defense_experiments.py:579-646 for one image (batch 1).  `model` is a FunctionalResNet50."""
    pred_clean = model(attacks_ref.normalize(x, mean, std)).argmax(dim=1)
    if attack_name == "fgsm":
        x_adv = attacks_ref.fgsm(model, x, y_true, eps=eps, mean=mean, std=std)
    elif attack_name == "pgd":
        x_adv = attacks_ref.pgd_linf(model, x, y_true, eps=eps, alpha=alpha, steps=steps, mean=mean, std=std,
                                     noise=noise)
    else:
        x_adv = attacks_ref.cw_l2(model, x, y_true, mean=mean, std=std, **(cw or {}))[0]
    pred_adv = model(attacks_ref.normalize(x_adv, mean, std)).argmax(dim=1)
    x_def = torch.from_numpy(defend_input(x_adv.numpy()))
    pred_def = model(attacks_ref.normalize(x_def, mean, std)).argmax(dim=1)
    s_adv = layer3_score(model.layer3_features(attacks_ref.normalize(x_adv, mean, std)))
    s_clean = layer3_score(model.layer3_features(attacks_ref.normalize(x, mean, std)))
    attack_success = int(pred_adv.item() != y_true.item())
    flag_adv = int(s_adv.item() > threshold)
    return {
        "clean_correct": bool(pred_clean.item() == y_true.item()),
        "attack_success": attack_success,
        "defense_preproc_success": int(pred_def.item() == y_true.item()),
        "detector_flags_clean": int(s_clean.item() > threshold),
        "detector_flags_adv": flag_adv,
        "detector_attack_success": int(attack_success == 1 and flag_adv == 0),
    }, {"x_adv": x_adv, "score_clean": float(s_clean), "score_adv": float(s_adv)}
