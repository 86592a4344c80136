"""Generates tests/golden/jpeg_cases.npz by running the REFERENCE's own `_jpeg_compress_batch` / `defend_input`
(defense_experiments.py:188-217, on this container's PIL / libjpeg-turbo) — build container only, the outputs are
committed.  Outputs are stored as the uint8 the decoder produced (the function returns them / 255).

    python oracle/gen_golden_jpeg.py
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.dont_write_bytecode = True

from oracle.gen_golden import import_reference  # noqa: E402

OUT = ROOT / "tests" / "golden"


def smooth_img(g, n, h, w, noise=0.05):
    yy, xx = torch.meshgrid(torch.linspace(0, 1, h), torch.linspace(0, 1, w), indexing="ij")
    base = torch.stack([0.5 + 0.5 * torch.sin(6 * xx + 3 * yy), yy * xx, 0.3 + 0.6 * xx])
    return base[None] + noise * torch.randn(n, 3, h, w, generator=g)


def cases():
    """name -> (float32 input [N,3,H,W], quality).  Values outside [0,1] are part of the contract (clamped)."""
    g = torch.Generator().manual_seed(2024)
    big = (smooth_img(g, 1, 224, 224, 0.08).clamp(0, 1) * 255).round().to(torch.uint8)
    out = {
        "img224": (big.float() / 255, 75),                       # stored as uint8, rebuilt as k/255
        "odd37x51": (smooth_img(g, 1, 37, 51) * 1.2 - 0.1, 75),
        "rand17x16": (torch.rand(1, 3, 17, 16, generator=g), 75),
        "one_pixel": (torch.rand(1, 3, 1, 1, generator=g), 75),
        "q30_64x48": (smooth_img(g, 1, 64, 48), 30),
        "q95_40x40": (torch.rand(2, 3, 40, 40, generator=g), 95),
        "saturated32": ((torch.rand(1, 3, 32, 32, generator=g) > 0.5).float(), 75),
        "q50_30x23": (smooth_img(g, 1, 30, 23), 50),
    }
    return out, big


def main():
    _, ref_def = import_reference()
    cs, big = cases()
    blob = {"img224_u8": big.numpy()}
    for name, (x, q) in cs.items():
        y = ref_def._jpeg_compress_batch(x, q)
        u8 = (y * 255).round().to(torch.uint8)
        assert torch.equal(u8.float() / 255, y)
        if name != "img224":
            blob[name + "_x"] = x.numpy()
        blob[name + "_q"] = np.int64(q)
        blob[name + "_out_u8"] = u8.numpy()
    # the combined defense with the JPEG switch on (defense_experiments.py:210-217, :173-174)
    g = torch.Generator().manual_seed(77)
    d = smooth_img(g, 2, 24, 40, 0.1) * 1.1 - 0.05
    ref_def._DEFENSE_USE_JPEG, ref_def._DEFENSE_JPEG_QUALITY = True, 75
    try:
        y = ref_def.defend_input(d)
    finally:
        ref_def._DEFENSE_USE_JPEG = False
    blob["defend_x"] = d.numpy()
    blob["defend_jpeg_out"] = y.numpy()
    np.savez_compressed(OUT / "jpeg_cases.npz", **blob)
    print("wrote", OUT / "jpeg_cases.npz", (OUT / "jpeg_cases.npz").stat().st_size, "bytes;", sorted(blob))


if __name__ == "__main__":
    main()
