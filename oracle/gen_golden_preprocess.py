"""Generates tests/golden/preprocess_cases.npz by running the REFERENCE's own transform
(`get_transform()`, defense_experiments.py:657-662 = Resize(256) -> CenterCrop(224) -> ToTensor on PIL images, this
container's torchvision + Pillow) — build container only; the outputs are committed.  Inputs are small recipes
(seeded noise / smooth gradients) rebuilt by `make_image`, except the reference's own picture/example.jpg, whose
decoded pixels are stored downsized-free as the uint8 crop result only (the decoded original is 2.3 MB).

    python oracle/gen_golden_preprocess.py
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.dont_write_bytecode = True

from oracle.gen_golden import import_reference  # noqa: E402

OUT = ROOT / "tests" / "golden"

SHAPES = [(256, 256), (300, 400), (683, 512), (500, 333), (240, 250), (224, 224), (256, 341), (1200, 1801), (257, 256),
          (255, 600)]


def make_image(h: int, w: int) -> np.ndarray:
    """Deterministic uint8 [h,w,3]: smooth structure + seeded noise (odd h+w) or pure seeded noise (even)."""
    rng = np.random.default_rng(h * 10007 + w)
    if (h + w) % 2 == 0:
        return rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    yy, xx = np.mgrid[0:h, 0:w]
    base = np.stack([128 + 100 * np.sin(xx / 17.0 + yy / 29.0), xx * 255.0 / w, yy * 255.0 / h], -1)
    return np.clip(base + rng.normal(0, 12, (h, w, 3)), 0, 255).astype(np.uint8)


def main():
    _, ref_def = import_reference()
    from PIL import Image
    tf = ref_def.get_transform()
    blob = {}
    for h, w in SHAPES:
        y = tf(Image.fromarray(make_image(h, w)))
        u8 = (y * 255).round().to(dtype=__import__("torch").uint8)
        assert (u8.float() / 255).equal(y)
        blob[f"out_{h}x{w}"] = u8.numpy()
    ex = Image.open("/root/reference/picture/example.jpg").convert("RGB")
    small = ex.crop((100, 80, 100 + 420, 80 + 330))          # a 330 x 420 window of the reference's example picture
    blob["example_crop_in"] = np.asarray(small)
    y = tf(small)
    blob["example_crop_out"] = (y * 255).round().numpy().astype(np.uint8)
    np.savez_compressed(OUT / "preprocess_cases.npz", **blob)
    print("wrote", (OUT / "preprocess_cases.npz").stat().st_size, "bytes", sorted(blob))


if __name__ == "__main__":
    main()
