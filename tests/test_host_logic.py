"""CPU-only checks of the boundary: the C-ABI library loads and exports every symbol the header
declares, the Python mirror keeps the reference's error behaviour, there is no CPU fallback, and
the multi-rank counter reduction works over gloo (world size 2)."""
import ctypes
import os
import re
import subprocess
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "advb200.h"


def _declared_symbols():
    text = HEADER.read_text()
    return sorted(set(re.findall(r"\b(advb200_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from advb200 import _lib
    lib = ctypes.CDLL(str(_lib.LIB_PATH))
    declared = _declared_symbols()
    assert len(declared) >= 18
    for name in declared:
        assert hasattr(lib, name), f"libadvb200.so does not export {name}"
    # the ctypes signature table covers exactly the header
    assert sorted(_lib.SIGNATURES) == declared
    assert _lib.load().advb200_version() >= 100


def test_argument_errors_do_not_need_a_gpu():
    from advb200 import _lib
    lib = _lib.load()
    rc = lib.advb200_quantize_f32(None, None, 16, 1, None)
    assert rc == -1 and b"quantize" in lib.advb200_last_error()
    with pytest.raises(ValueError):
        _lib.check(rc, "advb200_quantize_f32")
    assert lib.advb200_linf_step_f32(None, None, None, 0.1, 0.1, -1, 1, None) == -1
    assert lib.advb200_linf_step_f32(None, None, None, 0.1, 0.1, 4, 1, None) == -1
    assert lib.advb200_linf_step_f32(None, None, None, 0.1, 0.1, 0, 1, None) == 0      # empty batch is fine
    assert lib.advb200_cw_parts(150528) == 49


def test_reference_error_behaviour_is_kept():
    import advb200
    x = torch.zeros(1, 3, 224, 224)
    with pytest.raises(ValueError, match="mean/std must be 1D"):
        advb200.normalize_batch(x, torch.zeros(1, 3), torch.ones(3))
    with pytest.raises(ValueError, match="4D tensor"):
        advb200.normalize_batch(x[0], torch.zeros(3), torch.ones(3))
    with pytest.raises(ValueError, match="y_target must be provided"):
        advb200.cw_l2_attack(object(), x, torch.zeros(1, dtype=torch.long), mean=torch.zeros(3), std=torch.ones(3),
                             targeted=True)
    out = advb200.normalize_batch(torch.full((1, 3, 2, 2), 0.5), torch.tensor([0.5, 0.25, 0.0]),
                                  torch.tensor([0.5, 0.5, 0.25]))
    assert torch.equal(out[0, :, 0, 0], torch.tensor([0.0, 0.5, 2.0]))


def test_no_cpu_fallback():
    import advb200
    from oracle import weights
    if torch.cuda.is_available():
        pytest.skip("checks the no-GPU behaviour")
    model = weights.seeded_resnet50(0)
    x = torch.rand(1, 3, 224, 224)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        advb200.fgsm_attack(model, x, torch.zeros(1, dtype=torch.long), eps=0.03, mean=torch.zeros(3),
                            std=torch.ones(3))
    with pytest.raises(TypeError):
        advb200.native_model(torch.nn.Linear(4, 4))
    with pytest.raises(ValueError, match="CUDA tensor"):
        advb200.defend_input(x)


def test_transfer_targets_and_label_checks_without_a_gpu():
    """The transformer transfer targets have no CPU fallback either; label and noise-shape errors are raised by the host
    mirror before anything reaches the device."""
    if torch.cuda.is_available():
        pytest.skip("checks the no-GPU behaviour")
    import torchvision
    from advb200 import attack as A
    from advb200.transformers import NativeSwinT, NativeViTB16, _pad64
    vit = torchvision.models.vit_b_16(weights=None).eval()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        NativeViTB16(vit)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        NativeSwinT(torchvision.models.swin_t(weights=None).eval())
    assert [_pad64(c) for c in (48, 96, 288, 768, 1536)] == [64, 128, 320, 768, 1536]
    with pytest.raises(ValueError, match="labels must lie in"):
        A._check_labels(torch.tensor([0, 1000]), 2)
    with pytest.raises(ValueError, match="shape"):
        A._check_labels(torch.tensor([0, 1, 2]), 2)
    assert A._check_labels(torch.tensor([0, 999]), 2).tolist() == [0, 999]


def test_bench_reference_arm_and_shard_seeding():
    """bench.py: the strong-scaling shards of --config shard8192 tile the 8192 images for every rank count, and the
    reference arm reports which implementation it times."""
    import importlib.util
    from advb200 import dist as adist
    spec = importlib.util.spec_from_file_location("bench_mod", ROOT / "bench.py")
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    assert bench.REF_KIND in ("reference", "port")
    assert (bench.REF_KIND == "reference") == (ROOT / "oracle" / "_ref" / "attack.py").is_file()
    for world in (1, 2, 4, 8):
        spans = [adist.shard_range(8192, r, world) for r in range(world)]
        assert all((hi - lo) % 256 == 0 and lo % 256 == 0 for lo, hi in spans) and spans[-1][1] == 8192


def test_product_never_imports_the_oracle():
    pkg = ROOT / "image-recognition-adversarial-example-attack_b200"
    for p in pkg.rglob("*.py"):
        assert not re.search(r"^\s*(from|import)\s+oracle\b", p.read_text(), re.M), p


def test_threshold_tail_matches_oracle():
    import advb200
    from oracle import defenses_ref
    g = torch.Generator().manual_seed(3)
    for scale in (1.0, 0.01, 80.0):
        s = torch.rand(37, generator=g) * scale + 0.5 * scale
        assert advb200.calibrate_threshold_from_scores(s, 0.95) == defenses_ref.calibrate_threshold(s, 0.95)
    assert advb200.calibrate_threshold_from_scores(torch.empty(0)) == 5.0


def test_shard_ranges_cover_the_batch():
    from advb200 import dist as adist
    for n in (0, 1, 7, 8192, 4096 + 3):
        for world in (1, 2, 3, 8):
            spans = [adist.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(hi - lo for lo, hi in spans) - min(hi - lo for lo, hi in spans) <= 1


_WORKER = r"""
import os, sys, torch
sys.path.insert(0, sys.argv[1])
from advb200 import dist as adist
rank, world, _ = adist.init("gloo")
lo, hi = adist.shard_range(10, rank, world)
c = torch.tensor([hi - lo, rank + 1, 0, 0, 0, 0, hi - lo], dtype=torch.int64)
adist.allreduce_counters(c)
thr = adist.broadcast_scalar(25.5 if rank == 0 else -1.0)
mx = adist.max_over_ranks(float(rank))
adist.barrier()
assert c.tolist() == [10, 3, 0, 0, 0, 0, 10], c
assert thr == 25.5 and mx == 1.0
print("rank", rank, "ok")
"""


def test_counter_allreduce_world_size_2_gloo(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(_WORKER)
    pkg = str(ROOT / "image-recognition-adversarial-example-attack_b200")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29611", WORLD_SIZE="2")
    procs = [subprocess.Popen([sys.executable, str(script), pkg], env=dict(env, RANK=str(r), LOCAL_RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=120)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert all("ok" in o for o in outs)
