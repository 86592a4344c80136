"""Forward-only transfer target of the black-box evaluation (blackbox_transfer.py:25-54 loads VGG-19 / ViT-B/16 /
Swin-T; :233-251 classifies the ResNet-50 adversarials with them).  VGG-19 is conv-only and runs on the same
tcgen05 conv kernel as the attack path; the attention models are not built.
No CPU fallback."""
from __future__ import annotations

import ctypes
import weakref
from ctypes import c_void_p

import torch

from . import _lib


def _vgg19_parts(module: torch.nn.Module):
    feats = getattr(module, "features", None)
    cls = getattr(module, "classifier", None)
    if feats is None or cls is None:
        raise TypeError("expected a torchvision VGG-19 (features / classifier)")
    convs = [m for m in feats if isinstance(m, torch.nn.Conv2d)]
    fcs = [m for m in cls if isinstance(m, torch.nn.Linear)]
    cfg = [(c.in_channels, c.out_channels) for c in convs]
    want = [(3, 64), (64, 64), (64, 128), (128, 128), (128, 256), (256, 256), (256, 256), (256, 256), (256, 512),
            (512, 512), (512, 512), (512, 512), (512, 512), (512, 512), (512, 512), (512, 512)]
    if cfg != want or any(isinstance(m, torch.nn.BatchNorm2d) for m in feats) or \
            [(f.in_features, f.out_features) for f in fcs] != [(25088, 4096), (4096, 4096), (4096, 1000)]:
        raise TypeError("expected a torchvision VGG-19 without batch norm (16 convs, 3 linear layers)")
    if any(c.kernel_size != (3, 3) or c.stride != (1, 1) or c.padding != (1, 1) or c.bias is None for c in convs):
        raise TypeError("unexpected VGG-19 conv geometry")
    return convs, fcs


class NativeVGG19:
    """Device-resident packed VGG-19 (eval mode, frozen weights), forward only."""

    def __init__(self, module: torch.nn.Module, device=None):
        if not torch.cuda.is_available():
            raise RuntimeError("advb200 needs a CUDA device (B200); there is no CPU fallback")
        if module.training:
            raise ValueError("the model must be in eval() mode (dropout off)")
        convs, fcs = _vgg19_parts(module)
        dev = torch.device(device) if device is not None else next(module.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("the model must be on a CUDA device: advb200 has no CPU fallback")
        if dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        self.device = dev
        self.lib = _lib.load()
        keep = []

        def to_dev(t):
            t = t.detach().to(device=dev, dtype=torch.float32).contiguous()
            keep.append(t)
            return t

        arr = lambda ts: (c_void_p * len(ts))(*[t.data_ptr() for t in ts])
        with torch.cuda.device(dev):
            cw, cb = [to_dev(c.weight) for c in convs], [to_dev(c.bias) for c in convs]
            fw, fb = [to_dev(f.weight) for f in fcs], [to_dev(f.bias) for f in fcs]
            handle = c_void_p()
            rc = self.lib.advb200_vgg19_create(arr(cw), arr(cb), arr(fw), arr(fb), _lib.stream_ptr(), ctypes.byref(handle))
            _lib.check(rc, "advb200_vgg19_create")
        self.handle = handle
        self._ws = {}
        self.max_chunk = 128
        self._finalizer = weakref.finalize(self, self.lib.advb200_vgg19_destroy, handle)

    def _workspace(self, n: int):
        ent = self._ws.get(n)
        if ent is None:
            nbytes = int(self.lib.advb200_vgg19_workspace_bytes(self.handle, n))
            self._ws.clear()
            raw = torch.empty(nbytes + 1024, dtype=torch.uint8, device=self.device)
            ent = (raw, (raw.data_ptr() + 1023) // 1024 * 1024, nbytes)
            self._ws[n] = ent
        return ent

    def forward(self, x: torch.Tensor, mean: torch.Tensor, std: torch.Tensor) -> torch.Tensor:
        """fp32 logits [N,1000] of VGG-19 on normalize_batch(x, mean, std) (the normalisation is fused)."""
        if x.ndim != 4:
            raise ValueError("x must be a 4D tensor with shape [N, C, H, W]")
        if tuple(x.shape[1:]) != (3, 224, 224):
            raise ValueError("advb200 handles [N,3,224,224] inputs (the reference's CenterCrop(224) pipeline)")
        if not x.is_cuda:
            raise ValueError("x must live on the CUDA device of the native model (no CPU fallback)")
        if mean.ndim != 1 or std.ndim != 1:
            raise ValueError("mean/std must be 1D tensors with shape [C]")
        x = x.detach().to(torch.float32).contiguous()
        mean = mean.detach().to(device=x.device, dtype=torch.float32).contiguous()
        std = std.detach().to(device=x.device, dtype=torch.float32).contiguous()
        n = x.shape[0]
        out = torch.empty((n, 1000), dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            for lo in range(0, n, self.max_chunk):
                hi = min(n, lo + self.max_chunk)
                _, ws, nbytes = self._workspace(hi - lo)
                rc = self.lib.advb200_vgg19_forward(self.handle, _lib.ptr(x[lo:hi]), _lib.ptr(mean), _lib.ptr(std), hi - lo,
                                                    _lib.ptr(out[lo:hi]), None, c_void_p(ws), nbytes, _lib.stream_ptr())
                _lib.check(rc, "advb200_vgg19_forward")
        return out

    def __call__(self, x_normalised: torch.Tensor) -> torch.Tensor:
        """`model(normalize_batch(x))` call shape of blackbox_transfer.py:241 (input already normalised)."""
        dev = x_normalised.device
        return self.forward(x_normalised, torch.zeros(3, device=dev), torch.ones(3, device=dev))


def transfer_counts(target, x: torch.Tensor, x_adv: torch.Tensor, mean, std):
    """(successes, total) under the reference's criterion (blackbox_transfer.py:233-251): the attack transfers to an
    image when the target's prediction on the adversarial differs from ITS OWN prediction on the clean image."""
    y_clean = target.forward(x, mean, std).argmax(dim=1)
    y_adv = target.forward(x_adv, mean, std).argmax(dim=1)
    return int((y_clean != y_adv).sum()), int(y_clean.numel())


def transfer_success_rate(target, x: torch.Tensor, x_adv: torch.Tensor, mean, std) -> float:
    s, t = transfer_counts(target, x, x_adv, mean, std)
    return s / max(t, 1)
