"""Native ResNet-50 handle: packs a torchvision-shaped ResNet-50 into libadvb200 and exposes the two
calls the attack loop needs — logits (+ detector score) and the input gradient.

Replaces `model(normalize_batch(x, mean, std))` (attack.py:30,57,111), the autograd input gradient
(attack.py:32,59,130) and the layer3 partial forward (defense_experiments.py:232-238).
"""
from __future__ import annotations

import ctypes
import os
import weakref
from ctypes import c_float, c_void_p
from typing import Optional

import torch

from . import _lib

_DEFAULT_PRECISION = os.environ.get("ADVB200_PRECISION", "bf16")
_MAX_CHUNK = int(os.environ.get("ADVB200_MAX_CHUNK", "256"))


def set_default_precision(p: str) -> None:
    global _DEFAULT_PRECISION
    if p not in ("bf16", "fp32"):
        raise ValueError("precision must be 'bf16' or 'fp32'")
    _DEFAULT_PRECISION = p


def get_default_precision() -> str:
    return _DEFAULT_PRECISION


def _conv_bn_pairs(module):
    """(conv, bn) pairs of a torchvision ResNet-50 in libadvb200's canonical order."""
    try:
        pairs = [(module.conv1, module.bn1)]
        counts = []
        for layer in (module.layer1, module.layer2, module.layer3, module.layer4):
            counts.append(len(layer))
            for blk in layer:
                pairs += [(blk.conv1, blk.bn1), (blk.conv2, blk.bn2), (blk.conv3, blk.bn3)]
                if blk.downsample is not None:
                    pairs.append((blk.downsample[0], blk.downsample[1]))
        fc = module.fc
    except AttributeError as e:  # not a ResNet-shaped module
        raise TypeError("advb200 accelerates torchvision ResNet-50 only (no generic-model fallback)") from e
    shapes_ok = (counts == [3, 4, 6, 3] and len(pairs) == _lib.NUM_CONVS
                 and tuple(module.conv1.weight.shape) == (64, 3, 7, 7)
                 and tuple(fc.weight.shape) == (1000, 2048)
                 and module.layer1[0].conv2.stride == (1, 1) and module.layer2[0].conv2.stride == (2, 2))
    if not shapes_ok:
        raise TypeError("advb200 accelerates torchvision ResNet-50 (v1.5, 1000 classes) only")
    return pairs, fc


class NativeResNet50:
    """Device-resident packed ResNet-50 (eval mode, frozen weights)."""

    def __init__(self, module: torch.nn.Module, precision: Optional[str] = None, device=None):
        if not torch.cuda.is_available():
            raise RuntimeError("advb200 needs a CUDA device (B200); there is no CPU fallback")
        self.lib = _lib.load()
        self.precision = precision or _DEFAULT_PRECISION
        if self.precision not in ("bf16", "fp32"):
            raise ValueError("precision must be 'bf16' or 'fp32'")
        if module.training:
            raise ValueError("the model must be in eval() mode (BatchNorm uses running statistics)")
        dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        if dev.type != "cuda":
            raise RuntimeError("advb200 needs a CUDA device (B200); there is no CPU fallback")
        if dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        self.device = dev
        pairs, fc = _conv_bn_pairs(module)
        keep = []

        def to_dev(t):
            t = t.detach().to(device=self.device, dtype=torch.float32).contiguous()
            keep.append(t)
            return t

        arr = lambda ts: (c_void_p * len(ts))(*[t.data_ptr() for t in ts])
        with torch.cuda.device(self.device):
            ws = [to_dev(c.weight) for c, _ in pairs]
            ga = [to_dev(b.weight) for _, b in pairs]
            be = [to_dev(b.bias) for _, b in pairs]
            mu = [to_dev(b.running_mean) for _, b in pairs]
            va = [to_dev(b.running_var) for _, b in pairs]
            eps = (c_float * len(pairs))(*[float(b.eps) for _, b in pairs])
            fcw, fcb = to_dev(fc.weight), to_dev(fc.bias)
            handle = c_void_p()
            rc = self.lib.advb200_model_create(arr(ws), arr(ga), arr(be), arr(mu), arr(va), eps,
                                               _lib.ptr(fcw), _lib.ptr(fcb),
                                               _lib.PREC_BF16 if self.precision == "bf16" else _lib.PREC_FP32,
                                               _lib.stream_ptr(), ctypes.byref(handle))
            _lib.check(rc, "advb200_model_create")
        self.handle = handle
        self._ws = {}   # n -> (raw tensor, aligned ptr, bytes)
        self._pgd_bufs = {}   # n -> persistent iterate / gradient / label buffers of pgd_steps (stable pointers for the graph)
        self._finalizer = weakref.finalize(self, self.lib.advb200_model_destroy, handle)
        self.max_chunk = _MAX_CHUNK
        self._gen = 0   # bumped by every native call: tells the autograd wrapper whether its forward state survives

    # -- workspace ---------------------------------------------------------------------------
    def _workspace(self, n: int):
        ent = self._ws.get(n)
        if ent is None:
            nbytes = int(self.lib.advb200_workspace_bytes(self.handle, n))
            if nbytes <= 0:
                raise RuntimeError("advb200_workspace_bytes failed")
            # keep at most two workspaces alive (full chunk + tail chunk) to bound memory
            while len(self._ws) >= 2:
                self._ws.pop(next(iter(self._ws)))
            raw = torch.empty(nbytes + 1024, dtype=torch.uint8, device=self.device)
            base = (raw.data_ptr() + 1023) // 1024 * 1024
            ent = (raw, base, nbytes)
            self._ws[n] = ent
        return ent

    @staticmethod
    def _vec3(t: torch.Tensor, device) -> torch.Tensor:
        if t.ndim != 1:
            raise ValueError("mean/std must be 1D tensors with shape [C]")
        if t.numel() != 3:
            raise ValueError("mean/std must have 3 entries for an RGB ResNet-50")
        return t.detach().to(device=device, dtype=torch.float32).contiguous()

    def _check_x(self, x: torch.Tensor) -> torch.Tensor:
        if x.ndim != 4:
            raise ValueError("x must be a 4D tensor with shape [N, C, H, W]")
        if tuple(x.shape[1:]) != (3, 224, 224):
            raise ValueError("advb200 handles [N,3,224,224] inputs (the reference's CenterCrop(224) pipeline)")
        if not x.is_cuda:
            raise ValueError("x must live on the CUDA device of the native model (no CPU fallback)")
        if x.device != self.device:
            raise ValueError(f"x is on {x.device} but the native model lives on {self.device}")
        return x.detach().to(dtype=torch.float32).contiguous()

    # -- forward -----------------------------------------------------------------------------
    def forward(self, x, mean, std, *, want_logits=True, want_preds=False, want_score=False,
                stop_at_layer3=False):
        """Returns dict with any of 'logits' [N,1000], 'preds' [N] int64, 'score' [N]."""
        x = self._check_x(x)
        mean, std = self._vec3(mean, x.device), self._vec3(std, x.device)
        n = x.shape[0]
        self._gen += 1
        out = {}
        if want_logits and not stop_at_layer3:
            out["logits"] = torch.empty((n, 1000), dtype=torch.float32, device=x.device)
        if want_preds and not stop_at_layer3:
            out["preds"] = torch.empty((n,), dtype=torch.int64, device=x.device)
        if want_score:
            out["score"] = torch.empty((n,), dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            for lo in range(0, n, self.max_chunk):
                hi = min(n, lo + self.max_chunk)
                _, ws, nbytes = self._workspace(hi - lo)
                sl = lambda k: _lib.ptr(out[k][lo:hi]) if k in out else None
                rc = self.lib.advb200_forward(self.handle, _lib.ptr(x[lo:hi]), _lib.ptr(mean), _lib.ptr(std),
                                              hi - lo, sl("logits"), sl("preds"), int(bool(stop_at_layer3)),
                                              sl("score"), c_void_p(ws), nbytes, _lib.stream_ptr())
                _lib.check(rc, "advb200_forward")
        return out

    # -- forward + input gradient ------------------------------------------------------------
    def input_grad(self, x, mean, std, y_cmp, *, loss="ce", c=1.0, kappa=0.0, inv_n_global=None,
                   grad_out=None, want_logits=False, want_preds=False, want_loss=False, want_success=False,
                   success_out=None, preds_out=None, loss_out=None, logits_out=None):
        x = self._check_x(x)
        mean, std = self._vec3(mean, x.device), self._vec3(std, x.device)
        n = x.shape[0]
        self._gen += 1
        y_cmp = y_cmp.detach().to(device=x.device, dtype=torch.int64).contiguous()
        if y_cmp.shape != (n,):
            raise ValueError("labels must have shape [N]")
        kind = {"ce": _lib.LOSS_CE_MEAN, "cw": _lib.LOSS_CW_UNTGT, "cw_targeted": _lib.LOSS_CW_TGT}[loss]
        if inv_n_global is None:
            inv_n_global = 1.0 / n
        grad = grad_out if grad_out is not None else torch.empty_like(x)
        out = {"grad": grad}
        if want_logits or logits_out is not None:
            out["logits"] = logits_out if logits_out is not None else \
                torch.empty((n, 1000), dtype=torch.float32, device=x.device)
        if want_preds or preds_out is not None:
            out["preds"] = preds_out if preds_out is not None else \
                torch.empty((n,), dtype=torch.int64, device=x.device)
        if want_loss or loss_out is not None:
            out["loss"] = loss_out if loss_out is not None else \
                torch.empty((n,), dtype=torch.float32, device=x.device)
        if want_success or success_out is not None:
            out["success"] = success_out if success_out is not None else \
                torch.empty((n,), dtype=torch.uint8, device=x.device)
        with torch.cuda.device(x.device):
            for lo in range(0, n, self.max_chunk):
                hi = min(n, lo + self.max_chunk)
                _, ws, nbytes = self._workspace(hi - lo)
                sl = lambda k: _lib.ptr(out[k][lo:hi]) if k in out else None
                rc = self.lib.advb200_input_grad(self.handle, _lib.ptr(x[lo:hi]), _lib.ptr(mean), _lib.ptr(std),
                                                 _lib.ptr(y_cmp[lo:hi]), hi - lo, kind, float(c), float(kappa),
                                                 float(inv_n_global), sl("logits"), sl("preds"), sl("loss"),
                                                 sl("success"), _lib.ptr(grad[lo:hi]), c_void_p(ws), nbytes,
                                                 _lib.stream_ptr())
                _lib.check(rc, "advb200_input_grad")
        return out


    # -- backward for a caller-supplied dlogits ------------------------------------------------
    def backward(self, dlogits, std, n: int, grad_out=None):
        """Input gradient [n,3,224,224] for dlogits [n,1000], through the state the immediately preceding
        full `forward` of the same (single-chunk) batch left in the workspace (advb200_backward)."""
        if n > self.max_chunk:
            raise ValueError("backward works on one chunk: n <= max_chunk")
        dl = dlogits.detach().to(device=self.device, dtype=torch.float32).contiguous()
        if dl.shape != (n, 1000):
            raise ValueError("dlogits must have shape [N, 1000]")
        std = self._vec3(std, self.device)
        self._gen += 1
        grad = grad_out if grad_out is not None else torch.empty((n, 3, 224, 224), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            _, ws, nbytes = self._workspace(n)
            rc = self.lib.advb200_backward(self.handle, _lib.ptr(dl), _lib.ptr(std), n, _lib.ptr(grad), c_void_p(ws),
                                           nbytes, _lib.stream_ptr())
            _lib.check(rc, "advb200_backward")
        return grad

    # -- the PGD loop body, `steps` times, as one native call (CUDA-graph replay) -----------------
    def pgd_steps(self, x_adv, x_orig, mean, std, y_true, *, alpha, eps, steps, inv_n_global, use_graph=True):
        """x_adv <- clamp(project(x_adv + alpha*sign(grad CE(x_adv)))) `steps` times, in place (attack.py:55-63), for one
        chunk (n <= max_chunk).  The iteration runs on buffers this object owns (so the captured CUDA graph — ~105 kernel
        launches replayed as one host call per step — sees the same pointers on every attack); the chunk is copied in and the
        result copied back (two 0.6 MB/image device copies per attack)."""
        n = x_adv.shape[0]
        if n > self.max_chunk:
            raise ValueError("pgd_steps works on one chunk: n <= max_chunk")
        dev = self.device
        buf = self._pgd_bufs.get(n)
        if buf is None:
            while len(self._pgd_bufs) >= 2:
                self._pgd_bufs.pop(next(iter(self._pgd_bufs)))
            mk = lambda: torch.empty((n, 3, 224, 224), dtype=torch.float32, device=dev)
            buf = {"x_adv": mk(), "x_orig": mk(), "grad": mk(), "y": torch.empty((n,), dtype=torch.int64, device=dev),
                   "mean": torch.empty(3, dtype=torch.float32, device=dev), "std": torch.empty(3, dtype=torch.float32, device=dev)}
            self._pgd_bufs[n] = buf
        buf["x_adv"].copy_(x_adv)
        if x_orig is not None:
            buf["x_orig"].copy_(x_orig)
        buf["y"].copy_(y_true)
        buf["mean"].copy_(self._vec3(mean, dev))
        buf["std"].copy_(self._vec3(std, dev))
        self._gen += 1
        with torch.cuda.device(dev):
            _, ws, nbytes = self._workspace(n)
            rc = self.lib.advb200_pgd_steps(self.handle, _lib.ptr(buf["x_adv"]), _lib.ptr(buf["x_orig"]) if x_orig is not None else None,
                                            _lib.ptr(buf["mean"]), _lib.ptr(buf["std"]), _lib.ptr(buf["y"]), n, float(inv_n_global),
                                            float(alpha), float(eps), int(steps), _lib.ptr(buf["grad"]), c_void_p(ws), nbytes,
                                            int(bool(use_graph)), _lib.stream_ptr())
            _lib.check(rc, "advb200_pgd_steps")
        x_adv.copy_(buf["x_adv"])
        return x_adv

    # -- one fused attack step (bf16 mode) ----------------------------------------------------
    def attack_step(self, x_adv, x_orig, mean, std, y_true, *, alpha, eps, inv_n_global=None):
        """x_adv <- clamp(project(x_adv + alpha*sign(grad CE(x_adv)))) in place (attack.py:61-63; x_orig=None:
        no projection, attack.py:33-34).  The update runs in the epilogue of the stem's input-gradient kernel,
        so the gradient never goes through memory; bit-identical to input_grad + the linf step kernel."""
        if self.precision != "bf16":
            raise RuntimeError("attack_step is the bf16 fused path; use input_grad + the linf step in fp32 mode")
        if x_adv.dtype != torch.float32 or not x_adv.is_contiguous() or not x_adv.is_cuda:
            raise ValueError("x_adv must be a contiguous fp32 CUDA tensor (it is updated in place)")
        self._check_x(x_adv)
        self._gen += 1
        mean, std = self._vec3(mean, x_adv.device), self._vec3(std, x_adv.device)
        n = x_adv.shape[0]
        y = y_true.detach().to(device=x_adv.device, dtype=torch.int64).contiguous()
        if y.shape != (n,):
            raise ValueError("labels must have shape [N]")
        if x_orig is not None and (x_orig.shape != x_adv.shape or x_orig.dtype != torch.float32
                                   or not x_orig.is_contiguous()):
            raise ValueError("x_orig must match x_adv (contiguous fp32)")
        if inv_n_global is None:
            inv_n_global = 1.0 / n
        with torch.cuda.device(x_adv.device):
            for lo in range(0, n, self.max_chunk):
                hi = min(n, lo + self.max_chunk)
                _, ws, nbytes = self._workspace(hi - lo)
                rc = self.lib.advb200_attack_step(self.handle, _lib.ptr(x_adv[lo:hi]),
                                                  _lib.ptr(x_orig[lo:hi]) if x_orig is not None else None,
                                                  _lib.ptr(mean), _lib.ptr(std), _lib.ptr(y[lo:hi]), hi - lo,
                                                  float(inv_n_global), float(alpha), float(eps), None,
                                                  c_void_p(ws), nbytes, _lib.stream_ptr())
                _lib.check(rc, "advb200_attack_step")
        return x_adv


_CACHE: "weakref.WeakKeyDictionary[torch.nn.Module, dict]" = weakref.WeakKeyDictionary()


def native_model(model, precision: Optional[str] = None) -> NativeResNet50:
    """NativeResNet50 for `model` (an nn.Module or already a NativeResNet50), cached per module+precision.

    The packed copy snapshots the module's weights at first use; the attack path treats the
    classifier as frozen (no weight gradients)."""
    if isinstance(model, NativeResNet50):
        return model
    if not isinstance(model, torch.nn.Module):
        raise TypeError("model must be a torchvision ResNet-50 nn.Module or a NativeResNet50")
    precision = precision or _DEFAULT_PRECISION
    per = _CACHE.setdefault(model, {})
    nm = per.get(precision)
    if nm is None:
        _conv_bn_pairs(model)          # TypeError for anything that is not a ResNet-50
        try:
            dev = next(model.parameters()).device
        except StopIteration:
            raise TypeError("model has no parameters")
        if dev.type != "cuda":
            raise RuntimeError("the model must be on a CUDA device: advb200 has no CPU fallback")
        nm = NativeResNet50(model, precision=precision, device=dev)
        per[precision] = nm
    return nm


class _LogitsFn(torch.autograd.Function):
    """logits = native_forward(x) with x already normalised by the caller; backward = advb200_backward.  The
    workspace holds the state of ONE forward: if another native call ran in between, the forward of the chunk is
    run again before its backward (cheaper than keeping 5.7 GB of state per outstanding graph)."""

    @staticmethod
    def forward(ctx, x, owner):
        nm = owner.native
        xin = nm._check_x(x)
        ctx.owner = owner
        ctx.save_for_backward(xin)
        out = nm.forward(xin, owner._zero, owner._one)["logits"]
        ctx.gen = nm._gen
        ctx.in_dtype = x.dtype
        return out

    @staticmethod
    def backward(ctx, dlogits):
        owner = ctx.owner
        nm = owner.native
        (xin,) = ctx.saved_tensors
        n = xin.shape[0]
        grad = torch.empty_like(xin)
        dlogits = dlogits.contiguous()
        single = n <= nm.max_chunk
        for lo in range(0, n, nm.max_chunk):
            hi = min(n, lo + nm.max_chunk)
            if not (single and ctx.gen == nm._gen):
                nm.forward(xin[lo:hi], owner._zero, owner._one, want_logits=False, want_preds=True)
            nm.backward(dlogits[lo:hi], owner._one, hi - lo, grad_out=grad[lo:hi])   # bumps _gen: state consumed
        return grad.to(ctx.in_dtype), None


class NativeResNet50Module(torch.nn.Module):
    """`nn.Module` face of the native model for UNMODIFIED callers that write
    `logits = model(normalize_batch(x, mean, std)); g = torch.autograd.grad(loss, x)` themselves
    (visualized_attacks.py:337-358, defense_experiments.py:458-479).

    Takes the normalised batch [N,3,224,224] (the normalisation stays in the caller's autograd graph), returns
    fp32 logits [N,1000]; differentiable with respect to its input only (frozen weights).
    There is no CPU fallback: inputs must be CUDA tensors."""

    def __init__(self, model, precision: Optional[str] = None):
        super().__init__()
        self.native = native_model(model, precision)
        dev = self.native.device
        self._zero = torch.zeros(3, dtype=torch.float32, device=dev)
        self._one = torch.ones(3, dtype=torch.float32, device=dev)
        self.training = False

    def train(self, mode: bool = True):
        if mode:
            raise RuntimeError("the native model is an eval-mode, frozen-weight classifier")
        return self

    def forward(self, x_normalised: torch.Tensor) -> torch.Tensor:
        if x_normalised.ndim != 4:
            raise ValueError("x must be a 4D tensor with shape [N, C, H, W]")
        if torch.is_grad_enabled() and x_normalised.requires_grad:
            return _LogitsFn.apply(x_normalised, self)
        return self.native.forward(x_normalised, self._zero, self._one)["logits"]
