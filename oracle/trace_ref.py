"""Launch-by-launch oracle of the native ResNet-50 forward + input gradient (TEST INFRASTRUCTURE ONLY).

The network of attack.py:30-32 (`model(normalize_batch(x))`, `autograd.grad(loss, x)`) written out as explicit per-layer
forward and input-gradient formulas (no autograd), one named tensor per launch of the native plan
(image-recognition-adversarial-example-attack_b200/csrc/engine.cu: "fwd l2.0.c3+ds", "bwd l1.1.c2+c1", "p1", "g_top" ...).

Two uses:
  * free running (`walk(..., traced=None)`): every launch consumes the previous expected tensors — this must reproduce the
    autograd-based oracle (oracle/resnet50_bf16emu.py, itself pinned to torchvision) and is checked on the CPU;
  * teacher forced (`walk(..., traced=trace)`): every launch consumes the tensors the CUDA path ITSELF produced
    (advb200_debug_trace), so each launch of the real plan is compared with the oracle applied to its own inputs.  Only
    accumulation order separates the two sides then (the whole-network comparison cannot be tighter than bf16 noise:
    rounding to bf16 decorrelates under any perturbation), so the per-launch tolerance is one bf16 ulp.
"""
from __future__ import annotations

import re
from typing import Dict, Optional

import torch
import torch.nn.functional as F
from torch.nn.grad import conv2d_input

from .resnet50_bf16emu import _LAYERS, _MaxPoolSeqBf16, _rb, fold


def canonical(name: str) -> str:
    """'bwd l2.0.c2/*' and the per-class 'bwd l2.0.c1+ds/3' launches all write the tensor named 'bwd l2.0.c2' etc."""
    return re.sub(r"/[*0-9]$", "", name)


class ManualResNet50:
    def __init__(self, state_dict, precision: str = "bf16", device="cpu", dtype=torch.float64):
        self.bf16 = precision == "bf16"
        self.dtype, self.device = dtype, torch.device(device)
        sd = {k: v.detach().to(self.device) for k, v in state_dict.items()}
        self.w = {}
        self.blocks = []
        names = [("conv1", "bn1")]
        for li, nb in enumerate(_LAYERS, start=1):
            for b in range(nb):
                p = f"layer{li}.{b}"
                down = f"{p}.downsample.0.weight" in sd
                names += [(f"{p}.conv1", f"{p}.bn1"), (f"{p}.conv2", f"{p}.bn2"), (f"{p}.conv3", f"{p}.bn3")]
                if down:
                    names.append((f"{p}.downsample.0", f"{p}.downsample.1"))
                self.blocks.append((p, f"l{li}.{b}", 2 if (b == 0 and li > 1) else 1, down))
        for conv, bn in names:
            self.w[conv] = fold(sd, conv, bn, self.bf16, dtype)
        self.fc_w, self.fc_b = sd["fc.weight"].to(dtype), sd["fc.bias"].to(dtype)

    def R(self, t):
        return _rb(t) if self.bf16 else t

    def walk(self, x, mean, std, y, inv_n: float, traced: Optional[Dict[str, torch.Tensor]] = None,
             fused_tail: bool = True) -> Dict[str, torch.Tensor]:
        """All named tensors of one CE forward + input gradient.  `traced`: name -> NCHW tensor (teacher forcing)."""
        dt, dev = self.dtype, self.device
        exp: Dict[str, torch.Tensor] = {}

        def src(*names):           # the tensor a launch reads: the CUDA path's own if it was traced, else the expected one
            for nm in names:
                if traced is not None and nm in traced:
                    return traced[nm].to(device=dev, dtype=dt)
            for nm in names:
                if nm in exp:
                    return exp[nm]
            raise KeyError(names)

        R = self.R
        conv = lambda t, name, s=1, p=0: F.conv2d(t, self.w[name][0], stride=s, padding=p) + self.w[name][1]
        x = x.to(device=dev, dtype=dt)
        mean32, std32 = (t.to(device=dev, dtype=torch.float32) for t in (mean, std))
        xn = ((x.to(torch.float32) - mean32.view(1, -1, 1, 1)) / std32.view(1, -1, 1, 1)).to(dt)   # fp32, true division
        exp["fwd conv1"] = R(F.relu(conv(R(xn), "conv1", 2, 3)))
        c1 = src("fwd conv1")
        exp["p1"] = F.max_pool2d(c1, 3, 2, 1)
        h_name = "p1"
        fwd_names = []
        for p, lbl, stride, down in self.blocks:
            xin = src(h_name)
            a_n, b_n = f"fwd {lbl}.c1", f"fwd {lbl}.c2"
            exp[a_n] = R(F.relu(conv(xin, p + ".conv1")))
            exp[b_n] = R(F.relu(conv(src(a_n), p + ".conv2", stride, 1)))
            tail = fused_tail and self.bf16 and lbl.startswith("l1.") and not down
            if down:
                exp[f"fwd {lbl}.ds"] = R(conv(xin, p + ".downsample.0", stride))
                if traced is not None and f"fwd {lbl}.ds" in traced:          # un-fused plan: the branch was stored
                    idn = src(f"fwd {lbl}.ds")
                else:                                                          # fused conv3 + downsample: never rounded
                    idn = conv(xin, p + ".downsample.0", stride)
                out_n = f"fwd {lbl}.c3" if (traced is not None and f"fwd {lbl}.c3" in traced) else f"fwd {lbl}.c3+ds"
            else:
                idn = xin
                out_n = f"fwd {lbl}.c2+c3" if tail and not (traced is not None and f"fwd {lbl}.c3" in traced) else f"fwd {lbl}.c3"
            exp[out_n] = R(F.relu(conv(src(b_n), p + ".conv3") + idn))
            fwd_names.append((a_n, b_n, out_n, h_name))
            h_name = out_n
        l4 = src(h_name)
        exp["feat"] = l4.mean(dim=(2, 3)).view(l4.shape[0], -1, 1, 1)
        exp["logits"] = F.linear(src("feat").flatten(1), self.fc_w, self.fc_b).view(l4.shape[0], -1, 1, 1)
        z = src("logits").flatten(1)
        y = y.to(dev)
        exp["dlogits"] = ((torch.softmax(z, 1) - F.one_hot(y, z.shape[1]).to(dt)) * inv_n).view(z.shape[0], -1, 1, 1)
        exp["dfeat"] = (src("dlogits").flatten(1) @ self.fc_w).view(z.shape[0], -1, 1, 1)
        exp["g_top"] = R((l4 > 0).to(dt) * (src("dfeat") / 49.0))
        g_name = "g_top"
        for bi in range(len(self.blocks) - 1, -1, -1):
            p, lbl, stride, down = self.blocks[bi]
            a_n, b_n, out_n, in_n = fwd_names[bi]
            a, b, xin = src(a_n), src(b_n), src(in_n)
            g_sum = src(g_name)
            dgrad = lambda g, name, shape, s=1, pd=0: conv2d_input(shape, self.w[name][0], g, stride=s, padding=pd)
            gb_n, ga_n = f"bwd {lbl}.c3", f"bwd {lbl}.c2"
            exp[gb_n] = R((b > 0).to(dt) * dgrad(g_sum, p + ".conv3", b.shape))
            exp[ga_n] = R((a > 0).to(dt) * dgrad(src(gb_n), p + ".conv2", a.shape, stride, 1))
            gate = torch.ones_like(xin) if bi == 0 else (xin > 0).to(dt)      # p1's ReLU is handled by the max-pool backward
            g1 = dgrad(src(ga_n), p + ".conv1", xin.shape)
            if down:
                exp[f"bwd {lbl}.ds"] = R(dgrad(g_sum, p + ".downsample.0", xin.shape, stride, 0))
                if traced is not None and f"bwd {lbl}.ds" in traced:
                    skip, gx_n = src(f"bwd {lbl}.ds"), f"bwd {lbl}.c1"
                else:
                    skip, gx_n = dgrad(g_sum, p + ".downsample.0", xin.shape, stride, 0), f"bwd {lbl}.c1+ds"
            else:
                skip = g_sum
                tail = fused_tail and self.bf16 and lbl.startswith("l1.")
                gx_n = f"bwd {lbl}.c2+c1" if tail and not (traced is not None and f"bwd {lbl}.c1" in traced) else f"bwd {lbl}.c1"
            exp[gx_n] = R(gate * (g1 + skip))
            g_name = gx_n
        g_p1 = src(g_name)
        # max-pool backward: first-maximum routing, the (up to four) routed gradients added one by one in the storage
        # precision, then the stem ReLU gate
        if self.bf16:
            with torch.enable_grad():
                c1r = c1.detach().clone().requires_grad_(True)
                (routed,) = torch.autograd.grad(_MaxPoolSeqBf16.apply(c1r), c1r, g_p1)
        else:
            with torch.enable_grad():
                c1r = c1.detach().clone().requires_grad_(True)
                (routed,) = torch.autograd.grad(F.max_pool2d(c1r, 3, 2, 1), c1r, g_p1)
        exp["g_c1"] = (c1 > 0).to(dt) * routed
        gx = conv2d_input(tuple(x.shape), self.w["conv1"][0], src("g_c1"), stride=2, padding=3)
        exp["grad"] = (gx.to(torch.float32) / std32.view(1, -1, 1, 1)).to(dt) if self.bf16 else gx / std32.to(dt).view(1, -1, 1, 1)
        return exp


def compare(got: torch.Tensor, want: torch.Tensor, bf16: bool):
    """(relative L2 error, number of elements outside one storage ulp + a small absolute floor)."""
    got, want = got.double(), want.double()
    d = (got - want).abs()
    scale = float(want.abs().max()) or 1.0
    rel = float(d.norm() / want.norm()) if float(want.norm()) > 0 else float(d.norm())
    tol = (2.0 ** -7 if bf16 else 2.0 ** -17) * want.abs() + (2e-4 if bf16 else 2e-5) * scale
    return rel, int((d > tol).sum())
