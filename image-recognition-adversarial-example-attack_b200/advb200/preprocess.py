"""Image I/O mirror of the reference: `load_image` / `get_transform`
(defense_experiments.py:649-662; the same Compose at ResNet.py:27-31) and the ToPILImage of `--save_adv`
(ResNet.py:121-126).  File decoding / encoding stays with PIL on the host, as in the reference; the pixel work —
PIL's fixed-point antialiased resize, the centre crop, the /255 — runs on the device (advb200_preprocess_u8),
bit-exact with torchvision + PIL.  No CPU fallback."""
from __future__ import annotations

from pathlib import Path

import numpy as np
import torch

from . import _lib


class NativeTransform:
    """What `get_transform()` returns here: callable on a PIL image (or a uint8 HWC array / tensor), result
    float32 [3, crop, crop] on `device` — Resize(size) -> CenterCrop(crop) -> ToTensor."""

    def __init__(self, device, size: int = 256, crop: int = 224):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("advb200 preprocessing runs on a CUDA device (no CPU fallback)")
        self.size, self.crop = int(size), int(crop)

    def __call__(self, img) -> torch.Tensor:
        if isinstance(img, torch.Tensor):
            u8 = img
        else:
            if hasattr(img, "convert"):                       # PIL image
                img = np.asarray(img.convert("RGB"))
            u8 = torch.from_numpy(np.array(img, copy=True))          # (PIL / npz arrays can be read-only)
        if u8.dtype != torch.uint8 or u8.ndim != 3 or u8.shape[2] != 3:
            raise ValueError("expected a decoded RGB image: uint8 [H, W, 3]")
        u8 = u8.to(self.device).contiguous()
        h, w = int(u8.shape[0]), int(u8.shape[1])
        lib = _lib.load()
        out = torch.empty((3, self.crop, self.crop), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            nbytes = int(lib.advb200_preprocess_scratch_bytes(h, w, self.size, self.crop))
            if nbytes < 0:
                raise ValueError("image too small: the resized image would be smaller than the crop")
            scratch = torch.empty(nbytes + 256, dtype=torch.uint8, device=self.device)
            base = (scratch.data_ptr() + 255) // 256 * 256
            rc = lib.advb200_preprocess_u8(_lib.ptr(u8), h, w, self.size, self.crop, _lib.ptr(out),
                                           _lib.c_void_p(base), nbytes, _lib.stream_ptr())
        _lib.check(rc, "advb200_preprocess_u8")
        return out


def get_transform(device="cuda") -> NativeTransform:
    """defense_experiments.py:657-662."""
    return NativeTransform(device)


def load_image(path, device, transform=None) -> torch.Tensor:
    """defense_experiments.py:649-653: decoded by PIL on the host, preprocessed on the device -> [1,3,224,224]."""
    from PIL import Image
    tf = transform if transform is not None else NativeTransform(device)
    img = Image.open(Path(path)).convert("RGB")
    return tf(img).unsqueeze(0)


def to_pil_image(x: torch.Tensor):
    """transforms.ToPILImage()(x) for a float [3,H,W] CUDA tensor in [0,1] (ResNet.py:124): x*255 truncated to uint8
    on the device, then wrapped as a PIL image on the host."""
    from PIL import Image
    if x.ndim != 3 or x.shape[0] != 3 or not x.is_cuda:
        raise ValueError("expected a float [3, H, W] CUDA tensor")
    xc = x.detach().to(torch.float32).contiguous()
    h, w = int(xc.shape[1]), int(xc.shape[2])
    out = torch.empty((h, w, 3), dtype=torch.uint8, device=xc.device)
    with torch.cuda.device(xc.device):
        rc = _lib.load().advb200_to_u8_hwc(_lib.ptr(xc), _lib.ptr(out), h, w, _lib.stream_ptr())
    _lib.check(rc, "advb200_to_u8_hwc")
    return Image.fromarray(out.cpu().numpy(), mode="RGB")
