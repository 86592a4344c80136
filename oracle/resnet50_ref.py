"""Functional eval-mode ResNet-50 v1.5 from a torchvision state_dict (CPU, any float dtype).

Restates what `model(...)` does at attack.py:30,57,111 (torchvision.models.resnet50, third-party)
and the partial forward of defense_experiments.py:232-238, with plain torch.nn.functional calls.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

_LAYERS = (3, 4, 6, 3)


def _bn(sd, x, p):
    return F.batch_norm(x, sd[p + ".running_mean"], sd[p + ".running_var"], sd[p + ".weight"], sd[p + ".bias"],
                        False, 0.0, 1e-5)


def _bottleneck(sd, x, p, stride):
    out = F.relu(_bn(sd, F.conv2d(x, sd[p + ".conv1.weight"]), p + ".bn1"))
    out = F.relu(_bn(sd, F.conv2d(out, sd[p + ".conv2.weight"], stride=stride, padding=1), p + ".bn2"))
    out = _bn(sd, F.conv2d(out, sd[p + ".conv3.weight"]), p + ".bn3")
    if (p + ".downsample.0.weight") in sd:
        x = _bn(sd, F.conv2d(x, sd[p + ".downsample.0.weight"], stride=stride), p + ".downsample.1")
    return F.relu(out + x)


def forward(sd, x_norm: torch.Tensor, stop_at_layer3: bool = False):
    """x_norm: normalised [N,3,H,W].  Returns logits [N,1000] or the layer3 feature map."""
    x = F.relu(_bn(sd, F.conv2d(x_norm, sd["conv1.weight"], stride=2, padding=3), "bn1"))
    x = F.max_pool2d(x, 3, 2, 1)
    for li, nb in enumerate(_LAYERS, start=1):
        for b in range(nb):
            x = _bottleneck(sd, x, f"layer{li}.{b}", 2 if (b == 0 and li > 1) else 1)
        if li == 3 and stop_at_layer3:
            return x
    x = x.mean(dim=(2, 3))   # AdaptiveAvgPool2d(1) + flatten
    return F.linear(x, sd["fc.weight"], sd["fc.bias"])


class FunctionalResNet50:
    """Callable with the attributes the reference's detector probes for (conv1/layer3 are only
    checked with hasattr at defense_experiments.py:230)."""

    def __init__(self, state_dict, dtype=torch.float32):
        self.sd = {k: (v.to(dtype) if v.is_floating_point() else v) for k, v in state_dict.items()}

    def __call__(self, x_norm):
        return forward(self.sd, x_norm)

    def layer3_features(self, x_norm):
        return forward(self.sd, x_norm, stop_at_layer3=True)
