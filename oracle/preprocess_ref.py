"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's image preprocessing
(`transforms.Compose([Resize(256), CenterCrop(224), ToTensor()])`, ResNet.py:27-31, defense_experiments.py:657-662,
applied in `load_image` :649-653) for an already decoded uint8 RGB image, and of `--save_adv`'s ToPILImage
(ResNet.py:121-126).

The arithmetic lives in third-party code that is not vendored under /root/reference (no version pin in the
reference; here torchvision 0.26 + Pillow 12.2): torchvision computes the output size and the crop offsets, PIL's
`ImagingResample` (Resample.c) does the work — a separable convolution in FIXED POINT on uint8:
  * per output position a window [xmin, xmax) and bilinear ("triangle") weights widened by the scale factor when
    shrinking (antialiasing), normalised in double precision, then rounded to integers with 22 fractional bits
    (`precompute_coeffs`, `normalize_coeffs_8bpc`);
  * horizontal pass first, then vertical, each `clip8((sum(pixel*k) + 2^21) >> 22)` with a uint8 intermediate.
Pinned bit for bit against the reference's own transform on PIL by tests/golden/preprocess_cases.npz
(oracle/gen_golden_preprocess.py)."""
from __future__ import annotations

import math

import numpy as np

PRECISION_BITS = 32 - 8 - 2


def resized_size(h: int, w: int, size: int = 256):
    """torchvision.transforms.functional._compute_resized_output_size for an int `size` (shorter side -> size)."""
    short, long = (w, h) if w <= h else (h, w)
    new_short, new_long = size, int(size * long / short)
    return (new_long, new_short) if w <= h else (new_short, new_long)      # (new_h, new_w)


def crop_offsets(h: int, w: int, crop: int = 224):
    """torchvision center_crop: Python round() (half to even) of the half differences."""
    return int(round((h - crop) / 2.0)), int(round((w - crop) / 2.0))


def precompute_coeffs(in_size: int, out_size: int):
    """Resample.c precompute_coeffs + normalize_coeffs_8bpc for the bilinear filter over the whole axis.
    Returns (bounds [out,2] = (xmin, count), coeffs [out, ksize] int32)."""
    scale = filterscale = in_size / out_size
    if filterscale < 1.0:
        filterscale = 1.0
    support = 1.0 * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    bounds = np.zeros((out_size, 2), dtype=np.int32)
    kk = np.zeros((out_size, ksize), dtype=np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = 0.0 + (xx + 0.5) * scale
        xmin = int(center - support + 0.5)
        if xmin < 0:
            xmin = 0
        xmax = int(center + support + 0.5)
        if xmax > in_size:
            xmax = in_size
        xmax -= xmin
        w = []
        ww = 0.0
        for x in range(xmax):
            a = (x + xmin - center + 0.5) * ss
            a = -a if a < 0 else a
            v = 1.0 - a if a < 1.0 else 0.0
            w.append(v)
            ww += v
        for x in range(xmax):
            v = w[x] / ww if ww != 0.0 else w[x]
            kk[xx, x] = int(-0.5 + v * (1 << PRECISION_BITS)) if v < 0 else int(0.5 + v * (1 << PRECISION_BITS))
        bounds[xx] = (xmin, xmax)
    return bounds, kk


def _resample_axis0(img: np.ndarray, bounds, kk, lo: int, hi: int) -> np.ndarray:
    """Output rows [lo, hi) of the 1-D resampling of img [L, ...] uint8 along axis 0."""
    out = np.empty((hi - lo,) + img.shape[1:], dtype=np.uint8)
    src = img.astype(np.int64)
    for o in range(lo, hi):
        x0, n = int(bounds[o, 0]), int(bounds[o, 1])
        acc = np.tensordot(kk[o, :n].astype(np.int64), src[x0:x0 + n], axes=(0, 0)) + (1 << (PRECISION_BITS - 1))
        out[o - lo] = np.clip(acc >> PRECISION_BITS, 0, 255)
    return out


def resize_center_crop_u8(img: np.ndarray, size: int = 256, crop: int = 224) -> np.ndarray:
    """uint8 [H,W,3] -> uint8 [crop,crop,3]: Resize(size) (PIL bilinear, antialiased) then CenterCrop(crop).
    Only the cropped window is computed (the passes are separable, so this equals cropping the full result).
    Images whose resized side is smaller than the crop are outside the reference's use (CenterCrop would pad)."""
    h, w, _ = img.shape
    nh, nw = resized_size(h, w, size)
    if nh < crop or nw < crop:
        raise ValueError("resized image smaller than the crop")
    top, left = crop_offsets(nh, nw, crop)
    cur = img
    if nw != w:                                   # horizontal pass first (ImagingResample), only if needed
        hb, hk = precompute_coeffs(w, nw)
        cur = _resample_axis0(np.ascontiguousarray(cur.transpose(1, 0, 2)), hb, hk, left, left + crop).transpose(1, 0, 2)
    else:
        cur = cur[:, left:left + crop]
    if nh != h:
        vb, vk = precompute_coeffs(h, nh)
        cur = _resample_axis0(np.ascontiguousarray(cur), vb, vk, top, top + crop)
    else:
        cur = cur[top:top + crop]
    return np.ascontiguousarray(cur)


def load_image_tensor(img_u8: np.ndarray) -> np.ndarray:
    """The tensor `load_image` returns for a decoded RGB image: float32 [1,3,224,224] = uint8 / 255 (ToTensor)."""
    c = resize_center_crop_u8(img_u8)
    return (c.astype(np.float32) / np.float32(255.0)).transpose(2, 0, 1)[None]


def to_pil_u8(x: np.ndarray) -> np.ndarray:
    """ToPILImage on a float [3,H,W] tensor (ResNet.py:124): mul(255) then .byte() (truncation) -> uint8 [H,W,3]."""
    return (np.asarray(x, dtype=np.float32) * np.float32(255.0)).astype(np.float32).astype(np.uint8).transpose(1, 2, 0)
