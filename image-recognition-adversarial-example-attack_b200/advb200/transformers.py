"""Forward-only transformer transfer targets of the black-box evaluation: ViT-B/16 and Swin-T
(blackbox_transfer.py:36-54 `load_vit` / `load_swin`; :233-251 classify the ResNet-50 adversarial examples with them).

Composition in Python, arithmetic in libadvb200 (include/advb200.h, "Building blocks of the transformer transfer targets"):
every nn.Linear / patch-embedding conv runs on the tcgen05 GEMM kernel of the attack path (advb200_linear_forward), the
residual stream is fp32 with fused add + LayerNorm, attention is one CTA per (image, window, head).  Same module walk as
torchvision's `VisionTransformer.forward` / `SwinTransformer.forward` in eval mode (dropout and stochastic depth are
identities).  No CPU fallback: the module must live on a CUDA device."""
from __future__ import annotations

import ctypes
import weakref
from ctypes import c_void_p

import torch

from . import _lib


def _pad64(c: int) -> int:
    return (c + 63) // 64 * 64


class _Linear:
    """nn.Linear packed for the tensor-core kernel (weights bf16, bias fp32, channels padded to multiples of 64)."""

    def __init__(self, lib, weight: torch.Tensor, bias, dev):
        w = weight.detach().to(device=dev, dtype=torch.float32).contiguous()
        w = w.reshape(w.shape[0], -1)
        b = None if bias is None else bias.detach().to(device=dev, dtype=torch.float32).contiguous()
        self.cout, self.cin = w.shape
        self.cin_pad, self.cout_pad = _pad64(self.cin), _pad64(self.cout)
        self.lib = lib
        h = c_void_p()
        with torch.cuda.device(dev):
            rc = lib.advb200_linear_create(_lib.ptr(w), _lib.ptr(b), self.cin, self.cout, self.cin_pad, self.cout_pad,
                                           _lib.stream_ptr(), ctypes.byref(h))
        _lib.check(rc, "advb200_linear_create")
        self.handle = h
        self._fin = weakref.finalize(self, lib.advb200_linear_destroy, h)

    def __call__(self, x: torch.Tensor, y: torch.Tensor, m: int):
        rc = self.lib.advb200_linear_forward(self.handle, _lib.ptr(x), _lib.ptr(y), int(m), 0, None, _lib.stream_ptr())
        _lib.check(rc, "advb200_linear_forward")
        return y


def _f32(t, dev):
    return t.detach().to(device=dev, dtype=torch.float32).contiguous()


class _Base:
    def __init__(self, module: torch.nn.Module, device=None):
        if not torch.cuda.is_available():
            raise RuntimeError("advb200 needs a CUDA device (B200); there is no CPU fallback")
        if module.training:
            raise ValueError("the model must be in eval() mode (dropout / stochastic depth off)")
        dev = torch.device(device) if device is not None else next(module.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("the model must be on a CUDA device: advb200 has no CPU fallback")
        if dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        self.device = dev
        self.lib = _lib.load()
        self._bufs = {}

    # ---- thin wrappers over the C ABI (everything on torch's current stream) ----
    def _ln(self, x32, delta, norm, y, m, c_real, c_pad, mode):
        g, b, eps = norm
        rc = self.lib.advb200_add_layernorm(_lib.ptr(x32), _lib.ptr(delta), _lib.ptr(g), _lib.ptr(b), float(eps), _lib.ptr(y),
                                            int(m), int(c_real), int(c_pad), int(mode), _lib.stream_ptr())
        _lib.check(rc, "advb200_add_layernorm")

    def _gelu(self, x, n):
        _lib.check(self.lib.advb200_gelu_bf16(_lib.ptr(x), _lib.ptr(x), int(n), _lib.stream_ptr()), "advb200_gelu_bf16")

    def _attention(self, qkv, out, bias, n, T, heads, dim, ld_qkv, ld_out, c_real, gh, gw, ws, shift):
        rc = self.lib.advb200_window_attention_bf16(_lib.ptr(qkv), _lib.ptr(out), _lib.ptr(bias), int(n), int(T), int(heads), int(dim),
                                                    int(ld_qkv), int(ld_out), 0, int(c_real), int(2 * c_real), int(gh), int(gw), int(ws),
                                                    int(shift), float(dim) ** -0.5, _lib.stream_ptr())
        _lib.check(rc, "advb200_window_attention_bf16")

    def _head(self, x32, norm, n, tokens, t0, count, c, fc_w, fc_b):
        g, b, eps = norm
        feat = torch.empty((n, c), dtype=torch.float32, device=self.device)
        rc = self.lib.advb200_ln_pool_f32(_lib.ptr(x32), _lib.ptr(g), _lib.ptr(b), float(eps), _lib.ptr(feat), int(n), int(tokens),
                                          int(t0), int(count), int(c), int(c), _lib.stream_ptr())
        _lib.check(rc, "advb200_ln_pool_f32")
        classes = fc_w.shape[0]
        logits = torch.empty((n, classes), dtype=torch.float32, device=self.device)
        scratch = torch.empty(max(1, int(self.lib.advb200_fc_scratch_bytes(n, classes))), dtype=torch.uint8, device=self.device)
        rc = self.lib.advb200_fc_f32(_lib.ptr(feat), _lib.ptr(fc_w), _lib.ptr(fc_b), _lib.ptr(logits), _lib.ptr(scratch), int(n),
                                     int(classes), int(c), _lib.stream_ptr())
        _lib.check(rc, "advb200_fc_f32")
        return logits

    def _check(self, x, mean, std):
        if x.ndim != 4 or tuple(x.shape[1:]) != (3, 224, 224):
            raise ValueError("x must be [N,3,224,224] (the reference's CenterCrop(224) pipeline)")
        if not x.is_cuda or x.device != self.device:
            raise ValueError("x must live on the CUDA device of the native model (no CPU fallback)")
        if mean.ndim != 1 or std.ndim != 1:
            raise ValueError("mean/std must be 1D tensors with shape [C]")
        return x.detach().to(torch.float32).contiguous(), _f32(mean, self.device), _f32(std, self.device)

    def forward(self, x, mean, std, max_chunk: int = 128) -> torch.Tensor:
        """fp32 logits [N,1000] of model((x - mean) / std) (the normalisation is fused), in chunks of `max_chunk` images."""
        x, mean, std = self._check(x, mean, std)
        outs = []
        with torch.cuda.device(self.device):
            for lo in range(0, x.shape[0], max_chunk):
                outs.append(self._forward_chunk(x[lo:lo + max_chunk], mean, std))
        return torch.cat(outs) if outs else torch.empty((0, 1000), dtype=torch.float32, device=self.device)

    def __call__(self, x_normalised: torch.Tensor) -> torch.Tensor:
        """`model(normalize_batch(x))` call shape of blackbox_transfer.py:241 (input already normalised)."""
        dev = x_normalised.device
        return self.forward(x_normalised, torch.zeros(3, device=dev), torch.ones(3, device=dev))


class NativeViTB16(_Base):
    """torchvision `vit_b_16` (VisionTransformer: 16x16 patches, 12 layers, 12 heads, width 768), forward only."""

    def __init__(self, module: torch.nn.Module, device=None):
        super().__init__(module, device)
        try:
            layers = list(module.encoder.layers)
            ok = (module.patch_size == 16 and module.hidden_dim == 768 and module.image_size == 224 and len(layers) == 12
                  and tuple(module.conv_proj.weight.shape) == (768, 3, 16, 16) and layers[0].num_heads == 12
                  and tuple(module.heads.head.weight.shape) == (1000, 768))
        except AttributeError as e:
            raise TypeError("expected a torchvision ViT-B/16 (vit_b_16)") from e
        if not ok:
            raise TypeError("expected a torchvision ViT-B/16 (vit_b_16): 16x16 patches, 12 layers, width 768, 1000 classes")
        dev, lib = self.device, self.lib
        self.patch = _Linear(lib, module.conv_proj.weight, module.conv_proj.bias, dev)
        self.cls = _f32(module.class_token.reshape(-1), dev)
        self.pos = _f32(module.encoder.pos_embedding.reshape(197, 768), dev)
        norm = lambda ln: (_f32(ln.weight, dev), _f32(ln.bias, dev), ln.eps)
        self.layers = []
        for blk in layers:
            att = blk.self_attention
            self.layers.append({
                "ln1": norm(blk.ln_1), "ln2": norm(blk.ln_2),
                "qkv": _Linear(lib, att.in_proj_weight, att.in_proj_bias, dev),
                "proj": _Linear(lib, att.out_proj.weight, att.out_proj.bias, dev),
                "fc1": _Linear(lib, blk.mlp[0].weight, blk.mlp[0].bias, dev),
                "fc2": _Linear(lib, blk.mlp[3].weight, blk.mlp[3].bias, dev),
            })
        self.ln_f = norm(module.encoder.ln)
        self.fc_w, self.fc_b = _f32(module.heads.head.weight, dev), _f32(module.heads.head.bias, dev)

    def _buffers(self, n):
        b = self._bufs.get(n)
        if b is None:
            self._bufs.clear()
            dev, bf = self.device, torch.bfloat16
            m = n * 197
            b = {"cols": torch.empty((n * 196, 768), dtype=bf, device=dev), "patch": torch.empty((n * 196, 768), dtype=bf, device=dev),
                 "x32": torch.empty((m, 768), dtype=torch.float32, device=dev), "y": torch.empty((m, 768), dtype=bf, device=dev),
                 "qkv": torch.empty((m, 2304), dtype=bf, device=dev), "att": torch.empty((m, 768), dtype=bf, device=dev),
                 "delta": torch.empty((m, 768), dtype=bf, device=dev), "hid": torch.empty((m, 3072), dtype=bf, device=dev)}
            self._bufs[n] = b
        return b

    def _forward_chunk(self, x, mean, std):
        n = x.shape[0]
        m = n * 197
        b = self._buffers(n)
        lib = self.lib
        _lib.check(lib.advb200_patch_im2col_bf16(_lib.ptr(x), _lib.ptr(mean), _lib.ptr(std), _lib.ptr(b["cols"]), n, 224, 224, 16, 768,
                                                 _lib.stream_ptr()), "advb200_patch_im2col_bf16")
        self.patch(b["cols"], b["patch"], n * 196)                                     # conv_proj
        _lib.check(lib.advb200_vit_tokens_f32(_lib.ptr(b["patch"]), _lib.ptr(self.cls), _lib.ptr(self.pos), _lib.ptr(b["x32"]), n, 197,
                                              768, _lib.stream_ptr()), "advb200_vit_tokens_f32")    # class token + pos_embedding
        self._ln(b["x32"], None, self.layers[0]["ln1"], b["y"], m, 768, 768, 0)        # ln_1 of block 0
        for i, L in enumerate(self.layers):
            L["qkv"](b["y"], b["qkv"], m)
            self._attention(b["qkv"], b["att"], None, n, 197, 12, 64, 2304, 768, 768, 0, 0, 0, 0)
            L["proj"](b["att"], b["delta"], m)
            self._ln(b["x32"], b["delta"], L["ln2"], b["y"], m, 768, 768, 1)           # x = x + attn; y = ln_2(x)
            L["fc1"](b["y"], b["hid"], m)
            self._gelu(b["hid"], m * 3072)
            L["fc2"](b["hid"], b["delta"], m)
            nxt = self.layers[i + 1]["ln1"] if i + 1 < len(self.layers) else None
            if nxt is not None:
                self._ln(b["x32"], b["delta"], nxt, b["y"], m, 768, 768, 1)            # x = x + mlp; y = next ln_1(x)
            else:
                self._ln(b["x32"], b["delta"], self.ln_f, None, m, 768, 768, 1)        # x = x + mlp
        return self._head(b["x32"], self.ln_f, n, 197, 0, 1, 768, self.fc_w, self.fc_b)   # encoder.ln on the class token, head


class NativeSwinT(_Base):
    """torchvision `swin_t` (SwinTransformer v1: 4x4 patches, depths 2-2-6-2, widths 96-192-384-768, 7x7 windows)."""

    def __init__(self, module: torch.nn.Module, device=None):
        super().__init__(module, device)
        try:
            feats = list(module.features)
            ok = (len(feats) == 8 and tuple(feats[0][0].weight.shape) == (96, 3, 4, 4) and [len(feats[i]) for i in (1, 3, 5, 7)] == [2, 2, 6, 2]
                  and tuple(module.head.weight.shape) == (1000, 768) and list(feats[1][0].attn.window_size) == [7, 7]
                  and getattr(feats[1][0].attn, "logit_scale", None) is None)
        except (AttributeError, TypeError, IndexError) as e:
            raise TypeError("expected a torchvision Swin-T (swin_t)") from e
        if not ok:
            raise TypeError("expected a torchvision Swin-T (swin_t, version 1): depths 2-2-6-2, width 96, 7x7 windows")
        dev, lib = self.device, self.lib
        norm = lambda ln: (_f32(ln.weight, dev), _f32(ln.bias, dev), ln.eps)
        self.patch = _Linear(lib, feats[0][0].weight, feats[0][0].bias, dev)
        self.patch_norm = norm(feats[0][2])
        self.stages, self.merges = [], []
        grid = 56
        for si, fi in enumerate((1, 3, 5, 7)):
            blocks = []
            for blk in feats[fi]:
                att = blk.attn
                c = att.qkv.weight.shape[1]
                shift = int(att.shift_size[0]) if grid > 7 else 0            # no shift once the window covers the grid
                with torch.no_grad():
                    bias = att.get_relative_position_bias().reshape(att.num_heads, 49, 49)
                blocks.append({"c": c, "heads": att.num_heads, "shift": shift, "bias": _f32(bias, dev),
                               "ln1": norm(blk.norm1), "ln2": norm(blk.norm2),
                               "qkv": _Linear(lib, att.qkv.weight, att.qkv.bias, dev),
                               "proj": _Linear(lib, att.proj.weight, att.proj.bias, dev),
                               "fc1": _Linear(lib, blk.mlp[0].weight, blk.mlp[0].bias, dev),
                               "fc2": _Linear(lib, blk.mlp[3].weight, blk.mlp[3].bias, dev)})
            self.stages.append({"grid": grid, "c": blocks[0]["c"], "blocks": blocks})
            if si < 3:
                mg = feats[fi + 1]
                self.merges.append({"norm": norm(mg.norm), "red": _Linear(lib, mg.reduction.weight, None, dev)})
                grid //= 2
        self.ln_f = norm(module.norm)
        self.fc_w, self.fc_b = _f32(module.head.weight, dev), _f32(module.head.bias, dev)

    def _buffers(self, n):
        b = self._bufs.get(n)
        if b is None:
            self._bufs.clear()
            dev, bf = self.device, torch.bfloat16
            b = {"cols": torch.empty((n * 3136, 64), dtype=bf, device=dev), "stages": []}
            for st in self.stages:
                m, c = n * st["grid"] ** 2, st["c"]
                cp = _pad64(c)
                b["stages"].append({
                    "x32": torch.empty((m, cp), dtype=torch.float32, device=dev), "y": torch.empty((m, cp), dtype=bf, device=dev),
                    "qkv": torch.empty((m, _pad64(3 * c)), dtype=bf, device=dev),
                    "att": torch.zeros((m, cp), dtype=bf, device=dev),          # (padded columns stay zero: the kernel writes real ones)
                    "delta": torch.empty((m, cp), dtype=bf, device=dev), "hid": torch.empty((m, 4 * c), dtype=bf, device=dev),
                    "cat32": torch.empty((m // 4, 4 * c), dtype=torch.float32, device=dev),
                    "cat": torch.empty((m // 4, 4 * c), dtype=bf, device=dev)})
            self._bufs[n] = b
        return b

    def _forward_chunk(self, x, mean, std):
        n = x.shape[0]
        B = self._buffers(n)
        lib = self.lib
        _lib.check(lib.advb200_patch_im2col_bf16(_lib.ptr(x), _lib.ptr(mean), _lib.ptr(std), _lib.ptr(B["cols"]), n, 224, 224, 4, 64,
                                                 _lib.stream_ptr()), "advb200_patch_im2col_bf16")
        s0 = B["stages"][0]
        self.patch(B["cols"], s0["delta"], n * 3136)                                   # features[0][0]: 4x4/4 conv
        c0 = self.stages[0]["c"]
        # features[0][2]: LayerNorm(96) — its output IS the residual stream: x32 = LN(conv) via (set, normalise, set again)
        self._ln(s0["x32"], s0["delta"], self.patch_norm, s0["y"], n * 3136, c0, _pad64(c0), 2)
        self._ln(s0["x32"], s0["y"], self.stages[0]["blocks"][0]["ln1"], s0["y"], n * 3136, c0, _pad64(c0), 2)
        for si, st in enumerate(self.stages):
            b = B["stages"][si]
            g, c = st["grid"], st["c"]
            cp, m = _pad64(c), n * st["grid"] ** 2
            blocks = st["blocks"]
            for bi, L in enumerate(blocks):
                # here b["y"] holds norm1(x)
                L["qkv"](b["y"], b["qkv"], m)
                self._attention(b["qkv"], b["att"], L["bias"], n, 49, L["heads"], c // L["heads"], _pad64(3 * c), cp, c, g, g, 7, L["shift"])
                L["proj"](b["att"], b["delta"], m)
                self._ln(b["x32"], b["delta"], L["ln2"], b["y"], m, c, cp, 1)          # x = x + attn(norm1(x)); y = norm2(x)
                L["fc1"](b["y"], b["hid"], m)
                self._gelu(b["hid"], m * 4 * c)
                L["fc2"](b["hid"], b["delta"], m)
                if bi + 1 < len(blocks):
                    self._ln(b["x32"], b["delta"], blocks[bi + 1]["ln1"], b["y"], m, c, cp, 1)      # x = x + mlp; y = next norm1(x)
                else:
                    self._ln(b["x32"], b["delta"], L["ln2"], None, m, c, cp, 1)                      # x = x + mlp
            if si < 3:
                mg, nb = self.merges[si], B["stages"][si + 1]
                _lib.check(lib.advb200_patch_merge_gather_f32(_lib.ptr(b["x32"]), _lib.ptr(b["cat32"]), n, g, g, c, cp, _lib.stream_ptr()),
                           "advb200_patch_merge_gather_f32")
                self._ln(b["cat32"], None, mg["norm"], b["cat"], m // 4, 4 * c, 4 * c, 0)           # PatchMerging.norm
                mg["red"](b["cat"], nb["delta"], m // 4)                                           # PatchMerging.reduction
                nxt = self.stages[si + 1]
                self._ln(nb["x32"], nb["delta"], nxt["blocks"][0]["ln1"], nb["y"], m // 4, nxt["c"], _pad64(nxt["c"]), 2)
        last = B["stages"][3]
        return self._head(last["x32"], self.ln_f, n, 49, 0, 49, 768, self.fc_w, self.fc_b)    # norm, avgpool, head


def transfer_target(module: torch.nn.Module):
    """NativeVGG19 / NativeViTB16 / NativeSwinT for a torchvision module, by its class name."""
    name = type(module).__name__
    if name == "VisionTransformer":
        return NativeViTB16(module)
    if name == "SwinTransformer":
        return NativeSwinT(module)
    if name == "VGG":
        from .transfer import NativeVGG19
        return NativeVGG19(module)
    raise TypeError(f"no native transfer target for {name}")
