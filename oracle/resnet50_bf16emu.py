"""bf16-emulating restatement of the reference's ResNet-50 forward + input gradient (TEST INFRASTRUCTURE ONLY).

Same network as `model(normalize_batch(x))` at attack.py:30,57,111 and the same autograd input gradient as
attack.py:32,59,130 — but every tensor the bf16 product path stores in bf16 is rounded to bf16 at the same point here,
while all sums are carried in float64.  What remains between this emulator and the CUDA path is fp32 accumulation order
only, so the bf16 kernels (BN fold, epilogues, ReLU bit masks, fused launches, multi-class stride-2 input gradients,
max-pool routing) can be held to a tight tolerance, which a comparison against the un-rounded fp32 reference cannot do
(rounding activations to bf16 flips ReLU gates, and this network amplifies that).

Rounding points (image-recognition-adversarial-example-attack_b200/csrc):
  * folded weights  w * gamma/sqrt(var+eps)  -> bf16 (pack_kernel, pack_conv1_kernel); shift stays fp32
  * normalised image (x-mean)/std            -> bf16 (prep_hcol_kernel)
  * every conv epilogue output (after shift, residual / fused downsample branch, ReLU) -> bf16
  * every activation gradient (after the skip-gradient add and the ReLU gate) -> bf16; the max-pool backward adds its
    up-to-four routed gradients one by one in bf16 (window order), the stem input gradient stays fp32 (then / std)
  * avg-pool, fc, loss and its gradient: fp32 in the product, float64 here.
With `emulate=False` no rounding is applied and the functions are the plain reference network (pinned against
torchvision in tests/test_oracle_golden.py).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

_LAYERS = (3, 4, 6, 3)


def _rb(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.bfloat16).to(t.dtype)


class _RoundBoth(torch.autograd.Function):
    """bf16 rounding of an activation in the forward pass and of its gradient in the backward pass."""

    @staticmethod
    def forward(ctx, x):
        return _rb(x)

    @staticmethod
    def backward(ctx, g):
        return _rb(g)


class _RoundFwd(torch.autograd.Function):
    """bf16 rounding of the normalised image; its gradient (the stem input gradient) is fp32 in the product."""

    @staticmethod
    def forward(ctx, x):
        return _rb(x)

    @staticmethod
    def backward(ctx, g):
        return g


class _RoundBwd(torch.autograd.Function):
    """Identity forward (the value is a selection of bf16 values already); the gradient buffer is bf16."""

    @staticmethod
    def forward(ctx, x):
        return x.clone()

    @staticmethod
    def backward(ctx, g):
        return _rb(g)


class _MaxPoolSeqBf16(torch.autograd.Function):
    """3x3/2 max pool (pad 1, first maximum wins) whose backward adds the routed gradients one at a time in bf16, in
    ascending window order — maxpool_bwd_bf16x8_kernel."""

    @staticmethod
    def forward(ctx, x):
        out, idx = F.max_pool2d(x, 3, 2, 1, return_indices=True)
        ctx.save_for_backward(idx)
        ctx.in_shape = x.shape
        return out

    @staticmethod
    def backward(ctx, g):
        (idx,) = ctx.saved_tensors
        n, c, h, w = ctx.in_shape
        ho, wo = idx.shape[2], idx.shape[3]
        ih, iw = idx // w, idx % w
        oh = torch.arange(ho, device=g.device).view(1, 1, ho, 1)
        ow = torch.arange(wo, device=g.device).view(1, 1, 1, wo)
        di, dj = oh - ih // 2, ow - iw // 2            # each 0 or 1: which of the (up to) four windows of the input
        gin = torch.zeros(n, c, h * w, dtype=g.dtype, device=g.device)
        for i0 in (0, 1):
            for j0 in (0, 1):
                sel = ((di == i0) & (dj == j0)).to(g.dtype)
                gin = _rb(gin.scatter_add(2, idx.view(n, c, -1), (g * sel).view(n, c, -1)))
        return gin.view(n, c, h, w)


def fold(sd, conv: str, bn: str, emulate: bool, dtype):
    """BN folded into the conv: (weights * scale, shift), the scale computed in fp32 like bn_fold_kernel."""
    f32 = lambda k: sd[k].to(torch.float32)
    scale = f32(bn + ".weight") / torch.sqrt(f32(bn + ".running_var") + 1e-5)
    shift = f32(bn + ".bias") - f32(bn + ".running_mean") * scale
    w = f32(conv + ".weight") * scale.view(-1, 1, 1, 1)
    if emulate:
        w = _rb(w)
    return w.to(dtype), shift.to(dtype).view(1, -1, 1, 1)


class Bf16EmuResNet50:
    """Callable on the UN-normalised image batch; returns logits.  `layer3_features` gives the detector's feature map."""

    def __init__(self, state_dict, emulate: bool = True, dtype=torch.float64, device="cpu"):
        self.emulate, self.dtype, self.device = emulate, dtype, torch.device(device)
        sd = {k: v.detach().to(self.device) for k, v in state_dict.items()}
        self.w = {}
        names = [("conv1", "bn1")]
        for li, nb in enumerate(_LAYERS, start=1):
            for b in range(nb):
                p = f"layer{li}.{b}"
                names += [(f"{p}.conv1", f"{p}.bn1"), (f"{p}.conv2", f"{p}.bn2"), (f"{p}.conv3", f"{p}.bn3")]
                if f"{p}.downsample.0.weight" in sd:
                    names.append((f"{p}.downsample.0", f"{p}.downsample.1"))
        for conv, bn in names:
            self.w[conv] = fold(sd, conv, bn, emulate, dtype)
        self.fc_w, self.fc_b = sd["fc.weight"].to(dtype), sd["fc.bias"].to(dtype)
        self.R = _RoundBoth.apply if emulate else (lambda t: t)

    def _conv(self, x, name, stride=1, padding=0):
        w, shift = self.w[name]
        return F.conv2d(x, w, stride=stride, padding=padding) + shift

    def _block(self, x, p, stride):
        R = self.R
        a = R(F.relu(self._conv(x, p + ".conv1")))
        b = R(F.relu(self._conv(a, p + ".conv2", stride, 1)))
        out = self._conv(b, p + ".conv3")
        idn = self._conv(x, p + ".downsample.0", stride) if (p + ".downsample.0") in self.w else x
        return R(F.relu(out + idn))             # fused conv3 + downsample: the branch is never rounded on its own

    def features(self, x, mean, std, stop_at_layer3=False):
        x = x.to(device=self.device, dtype=self.dtype)
        mean, std = (t.to(device=self.device, dtype=torch.float32) for t in (mean, std))
        if self.emulate:     # the product normalises in fp32 with a true division, then rounds to bf16
            xn32 = (x.to(torch.float32) - mean.view(1, -1, 1, 1)) / std.view(1, -1, 1, 1)
            # keep the graph in `dtype`: same values as the fp32 expression up to one fp32 rounding before the bf16 one
            xn = (x - mean.to(self.dtype).view(1, -1, 1, 1)) / std.to(self.dtype).view(1, -1, 1, 1)
            xn = xn + (xn32.to(self.dtype) - xn).detach()
            xn = _RoundFwd.apply(xn)
        else:
            xn = (x - mean.to(self.dtype).view(1, -1, 1, 1)) / std.to(self.dtype).view(1, -1, 1, 1)
        h = self.R(F.relu(self._conv(xn, "conv1", 2, 3)))
        h = _RoundBwd.apply(_MaxPoolSeqBf16.apply(h)) if self.emulate else F.max_pool2d(h, 3, 2, 1)
        for li, nb in enumerate(_LAYERS, start=1):
            for b in range(nb):
                h = self._block(h, f"layer{li}.{b}", 2 if (b == 0 and li > 1) else 1)
            if li == 3 and stop_at_layer3:
                return h
        return h

    def __call__(self, x, mean, std):
        h = self.features(x, mean, std)
        return F.linear(h.mean(dim=(2, 3)), self.fc_w, self.fc_b)

    def layer3_features(self, x, mean, std):
        return self.features(x, mean, std, stop_at_layer3=True)

    def input_grad(self, x, y, mean, std, loss="ce", c=1.0, kappa=0.0, inv_n=None):
        """(gradient w.r.t. the un-normalised image, logits) for the CE mean (attack.py:31-32,58-59) or the CW margin
        objective without its L2 term (attack.py:114-127)."""
        xi = x.detach().to(device=self.device, dtype=self.dtype).clone().requires_grad_(True)
        y = y.to(self.device)
        logits = self(xi, mean, std)
        if loss == "ce":
            obj = F.cross_entropy(logits, y, reduction="sum") * (1.0 / x.shape[0] if inv_n is None else inv_n)
        else:
            onehot = F.one_hot(y, logits.shape[1]).to(logits.dtype)
            real = (logits * onehot).sum(1)
            other = (logits - 1e4 * onehot).amax(1)
            f = torch.clamp(other - real + kappa, min=0) if loss == "cw_targeted" else torch.clamp(real - other + kappa, min=0)
            obj = (c * f).sum()
        (g,) = torch.autograd.grad(obj, xi)
        return g, logits.detach()
