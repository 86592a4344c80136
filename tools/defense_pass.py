"""One warm launch of every HBM-bound defense kernel and of the CW-L2 update kernels at batch N inside a
cudaProfilerStart/Stop range (for `ncu --profile-from-start off`), all through the C ABI on preallocated buffers:
    python tools/defense_pass.py [N]"""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "image-recognition-adversarial-example-attack_b200")); sys.path.insert(0, str(ROOT))
import torch
from advb200 import _lib
from oracle import weights
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
lib = _lib.load()
P, S = _lib.ptr, _lib.stream_ptr
x = weights.synthetic_images(min(n, 32)).to(dev)
x = x.repeat((n + x.shape[0] - 1) // x.shape[0], 1, 1, 1)[:n].contiguous()
out = torch.empty_like(x)
planes, H, W, elems = n * 3, x.shape[2], x.shape[3], x.numel()
jbytes = int(lib.advb200_jpeg_scratch_bytes(n, H, W))
jscratch = torch.empty(jbytes, dtype=torch.uint8, device=dev)


per = x[0].numel()
parts = int(lib.advb200_cw_parts(per))
cw_w, cw_m, cw_v, cw_x0, cw_xa, cw_best, cw_g = (torch.empty_like(x) for _ in range(7))
cw_g.normal_(0, 1e-3)
cw_bl2 = torch.empty(n, device=dev)
cw_bs = torch.empty(n, dtype=torch.uint8, device=dev)
cw_succ = torch.zeros(n, dtype=torch.uint8, device=dev)
cw_pa = torch.empty(n * parts + (n + 3) // 4 + 1, device=dev)
cw_pb = torch.empty_like(cw_pa)
xa, x0 = x.clone(), x.clone()


def one():
    assert lib.advb200_linf_step_f32(P(xa), P(x0), P(cw_g), 2 / 255, 8 / 255, elems, 1, S()) == 0
    assert lib.advb200_cw_init_f32(P(x), P(cw_x0), P(cw_w), P(cw_m), P(cw_v), P(cw_xa), P(cw_best), P(cw_bl2), P(cw_bs), P(cw_pa),
                                   n, per, S()) == 0
    assert lib.advb200_cw_step_f32(P(cw_w), P(cw_m), P(cw_v), P(cw_x0), P(cw_g), P(cw_xa), P(cw_best), P(cw_bl2), P(cw_bs),
                                   P(cw_succ), P(cw_pa), P(cw_pb), None, n, per, 0.05, 0.0316, 0.9, 0.999, 1e-8, S()) == 0
    assert lib.advb200_smooth3x3_f32(P(x), P(out), planes, H, W, S()) == 0
    assert lib.advb200_median3x3_f32(P(x), P(out), planes, H, W, S()) == 0
    assert lib.advb200_quantize_f32(P(x), P(out), elems, 16, S()) == 0
    assert lib.advb200_defend_input_f32(P(x), P(out), planes, H, W, 16, S()) == 0
    assert lib.advb200_jpeg_roundtrip_f32(P(x), P(out), n, H, W, 75, P(jscratch), jbytes, S()) == 0


for _ in range(2):
    one()
torch.cuda.synchronize()
torch.cuda.profiler.start()
one()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
