"""Whole-network check that scheduling choices never change numerics: the plan-time autotuner (tile width incl. the
192-column tiles, CTA pairs, resident weights, staging slots) and the unit order of the multi-class launches only
reorder WHICH CTA computes WHAT; every accumulator sums its K blocks in the same order.  The same forward + input
gradient (bf16 product path, conditioned model, 16 images, one attacked away from a wrong label) runs in separate
processes — the switches are read once per process — and logits, loss and gradient must be equal bit for bit."""
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = Path(__file__).resolve().parent.parent

SCRIPT = r"""
import sys
sys.path.insert(0, sys.argv[2]); sys.path.insert(0, sys.argv[2] + "/image-recognition-adversarial-example-attack_b200")
import numpy as np, torch
import advb200
from oracle import weights
dev = torch.device("cuda:0")
nm = advb200.NativeResNet50(weights.conditioned_resnet50(0).to(dev), precision="bf16")
mean, std = weights.mean_std()
x = weights.mixed_images(16, seed=5).to(dev)
y = nm.forward(x, mean, std, want_preds=True)["preds"]
y[0] = (y[0] + 1) % 1000
for _ in range(2):          # the second pass runs on the tuned plan
    out = nm.input_grad(x, mean, std, y, loss="ce", want_logits=True)
np.savez(sys.argv[1], logits=out["logits"].cpu().numpy(), grad=out["grad"].cpu().numpy())
"""

VARIANTS = {"tuned": {}, "heuristic": {"ADVB200_NO_AUTOTUNE": "1"}, "classes_interleaved": {"ADVB200_CLS_ORDER": "minor"},
            "classes_major": {"ADVB200_CLS_ORDER": "major"}, "no_pairs": {"ADVB200_NO_AUTOTUNE": "1", "ADVB200_PAIR": "0"},
            # 192-column tiles (partial last column block) on every launch with >= 512 output channels
            "bn192": {"ADVB200_NO_AUTOTUNE": "1", "ADVB200_BN": "192"},
            "bn192_no_pairs": {"ADVB200_NO_AUTOTUNE": "1", "ADVB200_BN": "192", "ADVB200_PAIR": "0"}}


def test_plan_variants_are_bit_identical(tmp_path):
    got = {}
    for name, env in VARIANTS.items():
        f = tmp_path / f"{name}.npz"
        e = dict(os.environ); e.update(env)
        r = subprocess.run([sys.executable, "-c", SCRIPT, str(f), str(ROOT)], env=e, capture_output=True, text=True, timeout=900)
        assert r.returncode == 0, (name, r.stderr[-2000:])
        got[name] = np.load(f)
    base = got["tuned"]
    assert np.isfinite(base["grad"]).all() and float(np.abs(base["grad"]).max()) > 0
    for name, g in got.items():
        for k in ("logits", "grad"):
            assert np.array_equal(g[k], base[k]), (name, k, int((g[k] != base[k]).sum()))
