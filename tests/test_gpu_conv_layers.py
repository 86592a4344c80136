"""Layer-level parity of the tcgen05 conv kernel (every unique ResNet-50 conv geometry, forward
and input gradient, against torch fp32 convolutions on the same bf16-rounded operands), and the autotuner's
premise: every tile configuration it may pick gives bit-identical results."""
import sys
from pathlib import Path

import pytest
import torch
import torch.nn.functional as F

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "image-recognition-adversarial-example-attack_b200"))

pytestmark = pytest.mark.gpu

# n, h, w, cin, cout, k, stride — the 23 distinct conv geometries of torchvision's ResNet-50 after the stem
GEOMS = [
    (2, 56, 56, 64, 64, 1, 1), (2, 56, 56, 64, 64, 3, 1), (2, 56, 56, 64, 256, 1, 1), (2, 56, 56, 256, 64, 1, 1),
    (2, 56, 56, 256, 128, 1, 1), (2, 56, 56, 128, 128, 3, 2), (2, 56, 56, 256, 512, 1, 2),
    (3, 28, 28, 128, 512, 1, 1), (3, 28, 28, 512, 128, 1, 1), (3, 28, 28, 128, 128, 3, 1),
    (3, 28, 28, 512, 256, 1, 1), (3, 28, 28, 256, 256, 3, 2), (3, 28, 28, 512, 1024, 1, 2),
    (5, 14, 14, 256, 1024, 1, 1), (5, 14, 14, 1024, 256, 1, 1), (5, 14, 14, 256, 256, 3, 1),
    (5, 14, 14, 1024, 512, 1, 1), (5, 14, 14, 512, 512, 3, 2), (5, 14, 14, 1024, 2048, 1, 2),
    (5, 7, 7, 512, 2048, 1, 1), (5, 7, 7, 2048, 512, 1, 1), (5, 7, 7, 512, 512, 3, 1),
]
TOL = 1e-2          # bf16 outputs of fp32-accumulated sums of bf16 products vs the fp32 reference (relative L2)


@pytest.fixture(scope="module")
def env():
    import advb200  # noqa: F401
    from advb200 import _lib
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    return _lib, _lib.load(), torch.device("cuda:0")


def rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / (b.norm() + 1e-30))


def run_conv(env, x_nhwc, w, bias, add, mask, geom, relu, dgrad, out_shape):
    _lib, lib, dev = env
    n, h, wd, cin, cout, k, stride = geom
    out = torch.full(out_shape, float("nan"), dtype=torch.bfloat16, device=dev)
    rc = lib.advb200_test_conv(0, _lib.ptr(x_nhwc), _lib.ptr(w), _lib.ptr(bias), _lib.ptr(add), _lib.ptr(mask),
                               _lib.ptr(out), n, h, wd, cin, cout, k, stride, relu, dgrad, _lib.stream_ptr())
    _lib.check(rc, "advb200_test_conv")
    torch.cuda.synchronize()
    return out


def operands(geom, dev, seed):
    n, h, wd, cin, cout, k, stride = geom
    g = torch.Generator(device="cpu").manual_seed(seed)
    ho, wo = h // stride, wd // stride
    bq = lambda t: t.to(dev).to(torch.bfloat16).float()
    x = bq(torch.randn(n, cin, h, wd, generator=g))
    w = (torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).to(dev)
    b = torch.randn(cout, generator=g).to(dev)
    res = bq(torch.randn(n, cout, ho, wo, generator=g))
    gy = bq(torch.randn(n, cout, ho, wo, generator=g))
    act = bq(torch.randn(n, cin, h, wd, generator=g))
    addg = bq(torch.randn(n, cin, h, wd, generator=g))
    return x, w, b, res, gy, act, addg


def nhwc(t):
    return t.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16)


def forward_and_dgrad(env, geom, seed=1):
    n, h, wd, cin, cout, k, stride = geom
    ho, wo = h // stride, wd // stride
    x, w, b, res, gy, act, addg = operands(geom, env[2], seed)
    plain = k == 1 and stride == 2          # the downsample conv: no residual / ReLU forward, no add / mask backward
    no_add = plain or stride == 2           # the engine's stride-2 input gradients carry no residual operand
    out = run_conv(env, nhwc(x), w, b, None if plain else nhwc(res), None, geom, 0 if plain else 1, 0,
                   (n, ho, wo, cout))
    outg = run_conv(env, nhwc(gy), w, None, None if no_add else nhwc(addg), None if plain else nhwc(act), geom, 0, 1,
                    (n, h, wd, cin))
    return (x, w, b, res, gy, act, addg, plain, no_add), out, outg


@pytest.mark.parametrize("geom", GEOMS, ids=lambda g: "n%d_%dx%d_%dto%d_k%ds%d" % g)
def test_conv_forward_and_input_gradient_vs_torch_fp32(env, geom):
    env[1].advb200_test_force_tile(-1, -1, -1, -1)
    n, h, wd, cin, cout, k, stride = geom
    (x, w, b, res, gy, act, addg, plain, no_add), out, outg = forward_and_dgrad(env, geom)
    wq = w.to(torch.bfloat16).float()
    ref = F.conv2d(x, wq, b, stride=stride, padding=k // 2)
    if not plain:
        ref = F.relu(ref + res)
    assert not torch.isnan(out.float()).any() and not torch.isnan(outg.float()).any()
    assert rel(out.float().permute(0, 3, 1, 2), ref) < TOL
    refg = torch.nn.grad.conv2d_input((n, cin, h, wd), wq, gy, stride=stride, padding=k // 2)
    if not plain:
        refg = (refg if no_add else refg + addg) * (act > 0)
    assert rel(outg.float().permute(0, 3, 1, 2), refg) < TOL


CFG_GEOMS = [(2, 56, 56, 256, 64, 1, 1), (3, 28, 28, 128, 512, 1, 1), (3, 28, 28, 128, 128, 3, 1),
             (5, 14, 14, 512, 512, 3, 2), (5, 14, 14, 1024, 2048, 1, 2), (5, 7, 7, 512, 512, 3, 1)]


@pytest.mark.parametrize("geom", CFG_GEOMS, ids=lambda g: "n%d_%dx%d_%dto%d_k%ds%d" % g)
def test_every_autotuner_candidate_is_bit_identical(env, geom):
    """BN / CTA pair / resident weights / staging slots change the schedule, never the K order: same bits."""
    lib = env[1]
    lib.advb200_test_force_tile(-1, -1, -1, -1)
    _, base, baseg = forward_and_dgrad(env, geom)
    cout, cin = geom[4], geom[3]
    tried = 0
    bad = []
    try:
        for bn in (64, 128, 192, 256):
            for pair in (0, 1):
                for bres in (-1, 0):
                    for slots in (1, 2, 4):
                        # BN = 192 leaves the last column block partial (512 channels = 192 + 192 + 128): allowed from 256
                        # channels on (both the forward's cout and the input gradient's cin are GEMM column counts here)
                        if (min(cout, cin) < 256) if bn == 192 else (cout % bn or cin % bn):
                            continue
                        lib.advb200_test_force_tile(bn, pair, bres, slots)
                        _, out, outg = forward_and_dgrad(env, geom)
                        if not torch.equal(out, base):
                            bad.append((bn, pair, bres, slots, "forward", int((out != base).sum())))
                        if not torch.equal(outg, baseg):
                            bad.append((bn, pair, bres, slots, "dgrad", int((outg != baseg).sum())))
                        tried += 1
    finally:
        lib.advb200_test_force_tile(-1, -1, -1, -1)
    assert not bad, bad
    assert tried >= 12
