"""The allocation-free JPEG round trip, `advb200_jpeg_roundtrip_f32_noalloc` (include/advb200.h).

No counterpart in the reference: the entry point was added for callers that
cannot call cudaMalloc (CUDA graph capture, serving loops).  Its scratch is a static 24 MiB device array inside the
library, so the checks are: (1) through the C ABI it returns the SAME BITS as the existing caller-buffer entry point
`advb200_jpeg_roundtrip_f32` (integer / byte work: the bar is bit-exact, as in tests/test_gpu_jpeg.py), on calls made
back to back that leave stale bytes of another shape in the shared array; (2) sizes beyond the array are refused with a
clear message; (3) the call allocates nothing.  The file sorts last on purpose and
passes alone in a fresh process as well as after the other tests: it builds its own inputs and shares no fixtures.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import jpeg_ref  # noqa: E402

MAX_BYTES = 24 * 1024 * 1024           # ADVB200_JPEG_NOALLOC_MAX_BYTES


@pytest.fixture(scope="module")
def lib():
    from advb200 import _lib
    return _lib


def _images(shape, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(shape, generator=g) * 1.2 - 0.1              # includes values the clamp has to handle
    if shape[2] >= 16:
        x = torch.nn.functional.avg_pool2d(x, 5, 1, 2) * 1.5 - 0.2   # low-frequency content: AC terms survive
    return x.cuda().contiguous()


def _existing(lib, x, q):
    """advb200_jpeg_roundtrip_f32 with a caller-owned scratch buffer, through the C ABI."""
    L = lib.load()
    n, _, h, w = x.shape
    nbytes = int(L.advb200_jpeg_scratch_bytes(n, h, w))
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
    out = torch.full_like(x, float("nan"))
    lib.check(L.advb200_jpeg_roundtrip_f32(lib.ptr(x), lib.ptr(out), n, h, w, q, lib.ptr(scratch), nbytes, lib.stream_ptr()),
              "jpeg")
    return out


def _noalloc(lib, x, q, out=None):
    L = lib.load()
    n, _, h, w = x.shape
    out = torch.full_like(x, float("nan")) if out is None else out
    lib.check(L.advb200_jpeg_roundtrip_f32_noalloc(lib.ptr(x), lib.ptr(out), n, h, w, q, lib.stream_ptr()), "jpeg_noalloc")
    return out


def test_noalloc_equals_existing_entry_point_on_consecutive_calls(lib):
    # consecutive calls of different shapes and qualities: each one finds the previous call's decoded planes in the
    # shared array (other plane offsets, other row pitches); the last two repeat a shape, so a call also meets its own
    # stale bytes.  Odd sizes take the byte-store paths, 224 / 48 the 4-byte ones.
    cases = [((4, 3, 224, 224), 75), ((2, 3, 33, 517), 60), ((1, 3, 15, 9), 90), ((3, 3, 48, 300), 30),
             ((4, 3, 224, 224), 75), ((4, 3, 224, 224), 20)]
    for i, (shape, q) in enumerate(cases):
        x = _images(shape, seed=100 + i)
        got = _noalloc(lib, x, q)
        want = _existing(lib, x, q)
        torch.cuda.synchronize()
        assert not torch.isnan(got).any()
        assert torch.equal(got, want), (shape, q)               # bit-exact
    # and the pair of them against the CPU oracle (itself pinned bitwise to the reference's PIL round trip)
    x = _images((2, 3, 33, 517), seed=101)
    assert np.array_equal(_noalloc(lib, x, 60).cpu().numpy(), jpeg_ref.jpeg_compress_batch(x.cpu().numpy(), 60))


def test_noalloc_full_batch_and_in_place(lib):
    # the largest batch of the benchmarks fits: 256 x 224 x 224 needs 19.3 MB of the 24 MiB; x and out may alias
    x = _images((256, 3, 224, 224), seed=7)
    assert int(lib.load().advb200_jpeg_scratch_bytes(256, 224, 224)) <= MAX_BYTES
    want = _existing(lib, x, 75)
    got = _noalloc(lib, x, 75)
    assert torch.equal(got, want)
    xin = x.clone()
    _noalloc(lib, xin, 75, out=xin)
    assert torch.equal(xin, want)


def test_noalloc_rejects_sizes_beyond_the_array(lib):
    L = lib.load()
    x = torch.zeros(1, 3, 8, 8, device="cuda")
    out = torch.full_like(x, 7.0)
    for n, h, w in ((400, 224, 224), (1, 5000, 5000), (2 ** 20, 224, 224)):
        assert int(L.advb200_jpeg_scratch_bytes(n, h, w)) > MAX_BYTES
        # refused on the sizes alone, before any pointer is used: nothing is launched, `out` is untouched
        rc = L.advb200_jpeg_roundtrip_f32_noalloc(lib.ptr(x), lib.ptr(out), n, h, w, 75, lib.stream_ptr())
        assert rc == -1                                          # ADVB200_ERR_ARG
        msg = L.advb200_last_error().decode()
        assert "jpeg_noalloc" in msg and str(MAX_BYTES) in msg and "advb200_jpeg_roundtrip_f32" in msg, msg
    with pytest.raises(ValueError):
        lib.check(L.advb200_jpeg_roundtrip_f32_noalloc(lib.ptr(x), lib.ptr(out), 1, 8, 8, 0, lib.stream_ptr()), "jpeg_noalloc")
    with pytest.raises(ValueError):
        lib.check(L.advb200_jpeg_roundtrip_f32_noalloc(None, lib.ptr(out), 1, 8, 8, 75, lib.stream_ptr()), "jpeg_noalloc")
    torch.cuda.synchronize()
    assert bool((out == 7.0).all())
    # n_img == 0 is a no-op, as for the existing entry point
    assert L.advb200_jpeg_roundtrip_f32_noalloc(None, None, 0, 224, 224, 75, lib.stream_ptr()) == 0


def test_noalloc_call_allocates_nothing(lib):
    x = _images((8, 3, 224, 224), seed=3)
    out = torch.empty_like(x)
    _noalloc(lib, x, 75, out=out)                                # first use on this device: the one symbol lookup
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    stats0 = (torch.cuda.memory_allocated(), torch.cuda.memory_reserved())
    L = lib.load()
    for _ in range(20):
        assert L.advb200_jpeg_roundtrip_f32_noalloc(lib.ptr(x), lib.ptr(out), 8, 224, 224, 75, lib.stream_ptr()) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert (torch.cuda.memory_allocated(), torch.cuda.memory_reserved()) == stats0   # nothing through torch
    assert free1 >= free0                                        # nothing through the driver either
    # and it is capturable: a cudaMalloc inside a stream capture would invalidate the capture
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    out2 = torch.full_like(x, float("nan"))
    with torch.cuda.stream(s):
        g.capture_begin()
        rc = L.advb200_jpeg_roundtrip_f32_noalloc(lib.ptr(x), lib.ptr(out2), 8, 224, 224, 75,
                                                  ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        g.capture_end()
    assert rc == 0
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out2, out)
