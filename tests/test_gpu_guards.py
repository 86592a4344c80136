"""Own bounds checks (compute-sanitizer is closed on this GPU pool — profiles/r02_sanitizer_closed.txt): every caller-owned
buffer a C-ABI call writes is embedded between poisoned guard zones, and the guards must be untouched afterwards; every
byte of the output proper must have been written (NaN canary).  Whole-network passes (all conv launch kinds of the real
plan, both precisions) and the elementwise / CW / JPEG kernels, at sizes that leave partial tiles everywhere."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import weights  # noqa: E402

GUARD = 1 << 20          # bytes on each side


class Guarded:
    """A tensor view in the middle of a larger allocation whose margins are filled with a byte pattern."""

    def __init__(self, shape, dtype, dev, fill=None, align=1024):
        self.nbytes = int(torch.tensor([], dtype=dtype).element_size()) * int(torch.Size(shape).numel())
        self.raw = torch.full((2 * GUARD + self.nbytes + align,), 0xA5, dtype=torch.uint8, device=dev)
        off = GUARD + (-(self.raw.data_ptr() + GUARD)) % align
        self.off = off
        self.view = self.raw[off:off + self.nbytes].view(dtype).view(shape)
        if fill is not None:
            self.view.fill_(fill)

    def intact(self):
        lo, hi = self.raw[:self.off], self.raw[self.off + self.nbytes:]
        return bool((lo == 0xA5).all()) and bool((hi == 0xA5).all())


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
@pytest.mark.parametrize("n", [1, 3])
def test_whole_network_pass_stays_inside_its_workspace_and_outputs(precision, n):
    import ctypes
    import advb200
    from advb200 import _lib
    dev = torch.device("cuda:0")
    lib = _lib.load()
    nm = advb200.NativeResNet50(weights.conditioned_resnet50(0).to(dev), precision=precision)
    mean, std = (t.to(dev) for t in weights.mean_std())
    x = weights.mixed_images(n, seed=5).to(dev)
    y = torch.zeros(n, dtype=torch.int64, device=dev)
    need = int(lib.advb200_workspace_bytes(nm.handle, n))
    ws = Guarded((need,), torch.uint8, dev)
    grad = Guarded((n, 3, 224, 224), torch.float32, dev, fill=float("nan"), align=16)
    logits = Guarded((n, 1000), torch.float32, dev, fill=float("nan"), align=16)
    preds = Guarded((n,), torch.int64, dev, align=16)
    loss = Guarded((n,), torch.float32, dev, fill=float("nan"), align=16)
    succ = Guarded((n,), torch.uint8, dev, align=16)
    for _ in range(2):        # the first call also runs the plan-time autotuner on the same workspace
        rc = lib.advb200_input_grad(nm.handle, _lib.ptr(x), _lib.ptr(mean), _lib.ptr(std), _lib.ptr(y), n, _lib.LOSS_CE_MEAN, 1.0, 0.0,
                                    1.0 / n, _lib.ptr(logits.view), _lib.ptr(preds.view), _lib.ptr(loss.view), _lib.ptr(succ.view),
                                    _lib.ptr(grad.view), ctypes.c_void_p(ws.view.data_ptr()), need, _lib.stream_ptr())
        _lib.check(rc, "advb200_input_grad")
    score = Guarded((n,), torch.float32, dev, fill=float("nan"), align=16)
    _lib.check(lib.advb200_forward(nm.handle, _lib.ptr(x), _lib.ptr(mean), _lib.ptr(std), n, None, None, 1, _lib.ptr(score.view),
                                   ctypes.c_void_p(ws.view.data_ptr()), need, _lib.stream_ptr()), "advb200_forward")
    torch.cuda.synchronize()
    for name, gbuf in (("workspace", ws), ("grad", grad), ("logits", logits), ("preds", preds), ("loss", loss), ("success", succ),
                       ("score", score)):
        assert gbuf.intact(), f"{name}: guard zone overwritten"
    for name, gbuf in (("grad", grad), ("logits", logits), ("loss", loss), ("score", score)):
        assert not bool(torch.isnan(gbuf.view).any()), f"{name}: not every element was written"
    # a workspace one byte too small is refused, not overrun
    rc = lib.advb200_input_grad(nm.handle, _lib.ptr(x), _lib.ptr(mean), _lib.ptr(std), _lib.ptr(y), n, _lib.LOSS_CE_MEAN, 1.0, 0.0, 1.0,
                                None, None, None, None, _lib.ptr(grad.view), ctypes.c_void_p(ws.view.data_ptr() + 1024), need - 1024,
                                _lib.stream_ptr())
    assert rc == -1


@pytest.mark.parametrize("shape", [(3, 3, 224, 224), (2, 3, 37, 51), (1, 1, 5, 3)])
def test_elementwise_cw_and_jpeg_kernels_stay_inside_their_buffers(shape):
    from advb200 import _lib
    dev = torch.device("cuda:0")
    lib = _lib.load()
    P, S = _lib.ptr, _lib.stream_ptr
    n, c, h, w = shape
    x = torch.rand(shape, device=dev)
    elems = x.numel()
    bufs = {}

    def out(name, dtype=torch.float32, shp=shape, fill=float("nan")):
        bufs[name] = Guarded(shp, dtype, dev, fill=fill, align=16)
        return bufs[name].view

    g = torch.randn(shape, device=dev)
    xa = out("linf"); xa.copy_(x)
    assert lib.advb200_linf_step_f32(P(xa), P(x), P(g), 2 / 255, 8 / 255, elems, 1, S()) == 0
    noise = torch.empty(shape, device=dev).uniform_(-0.03, 0.03)
    assert lib.advb200_random_start_f32(P(out("start")), P(x), P(noise), elems, S()) == 0
    assert lib.advb200_smooth3x3_f32(P(x), P(out("smooth")), n * c, h, w, S()) == 0
    assert lib.advb200_median3x3_f32(P(x), P(out("median")), n * c, h, w, S()) == 0
    assert lib.advb200_quantize_f32(P(x), P(out("quant")), elems, 16, S()) == 0
    assert lib.advb200_defend_input_f32(P(x), P(out("defend")), n * c, h, w, 16, S()) == 0
    if c == 3:
        jb = int(lib.advb200_jpeg_scratch_bytes(n, h, w))
        js = Guarded((jb,), torch.uint8, dev, align=16)
        bufs["jpeg_scratch"] = js
        assert lib.advb200_jpeg_roundtrip_f32(P(x), P(out("jpeg")), n, h, w, 75, P(js.view), jb, S()) == 0
    per = c * h * w
    parts = int(lib.advb200_cw_parts(per))
    names = ("x0", "w", "m", "v", "x_adv", "best_adv")
    st = [out("cw_" + k) for k in names]
    bl2 = out("cw_best_l2", shp=(n,))
    bs = out("cw_best_success", dtype=torch.uint8, shp=(n,), fill=7)
    pa = out("cw_pa", shp=(n * parts + (n + 3) // 4 + 1,))
    pb = out("cw_pb", shp=(n * parts + (n + 3) // 4 + 1,))
    l2 = out("cw_l2", shp=(n,))
    assert lib.advb200_cw_init_f32(P(x), *[P(t) for t in st], P(bl2), P(bs), P(pa), n, per, S()) == 0
    succ = torch.ones(n, dtype=torch.uint8, device=dev)
    assert lib.advb200_cw_step_f32(P(st[1]), P(st[2]), P(st[3]), P(st[0]), P(g), P(st[4]), P(st[5]), P(bl2), P(bs), P(succ), P(pa), P(pb),
                                   P(l2), n, per, 0.05, 0.0316, 0.9, 0.999, 1e-8, S()) == 0
    assert lib.advb200_cw_finalize_f32(P(st[4]), P(st[5]), P(bs), P(out("cw_final")), n, per, S()) == 0
    torch.cuda.synchronize()
    for name, gbuf in bufs.items():
        assert gbuf.intact(), f"{name}: guard zone overwritten"
        if gbuf.view.dtype == torch.float32 and name not in ("cw_pa", "cw_pb"):
            assert not bool(torch.isnan(gbuf.view).any()), f"{name}: not every element was written"
