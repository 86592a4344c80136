"""Secondary measurements of SURVEY section 8(d), one JSON object on stdout:
  * HBM-bound update / defense kernels at N=256: achieved GB/s against MEASURED_PEAKS hbm_gbs
    (algorithmic bytes per element from DESIGN.md section 4.2);
  * config 3: CW-L2 (100 steps, c=1, lr=0.01), untargeted and targeted(805), batch 256, images/s;
  * config 5 (per GPU share): defense sweep FGSM+PGD x eps {4,8}/255 through the batched evaluator.
Usage: python tools/workloads_bench.py [batch] > gpurun_out/workloads.json"""
import ctypes
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "image-recognition-adversarial-example-attack_b200"))
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402
import advb200  # noqa: E402
from advb200 import _lib, defenses  # noqa: E402
from oracle import weights  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").is_file() else {}
PBW = float(peaks.get("hbm_gbs", 6650.0))
lib = _lib.load()


def timed(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


out = {"batch": n, "hbm_peak_gbs": PBW, "kernels": {}}
x = weights.synthetic_images(min(n, 64)).to(dev)
x = x.repeat((n + x.shape[0] - 1) // x.shape[0], 1, 1, 1)[:n].contiguous()
elems = x.numel()
g = torch.randn_like(x)
x0 = x.clone()
xa = x.clone()


def rec(name, ms, bytes_per_elem):
    gbs = bytes_per_elem * elems / (ms * 1e-3) / 1e9
    out["kernels"][name] = {"ms": ms, "bytes_per_elem": bytes_per_elem, "gbs": gbs, "frac_of_hbm_peak": gbs / PBW}


# every kernel is called through the C ABI on preallocated buffers (the Python wrappers' allocation + context-manager
# overhead is comparable to these 50-100 us kernels and would make the loop host-bound); each working set (> 300 MB)
# is larger than the 126 MB L2
from ctypes import c_void_p  # noqa: E402
P, S = _lib.ptr, _lib.stream_ptr
out_buf = torch.empty_like(x)
planes, H, W = n * 3, x.shape[2], x.shape[3]
rec("linf_step(project)", timed(lambda: lib.advb200_linf_step_f32(P(xa), P(x0), P(g), 2 / 255, 8 / 255, elems, 1, S())), 16)
rec("linf_step(fgsm)", timed(lambda: lib.advb200_linf_step_f32(P(xa), None, P(g), 8 / 255, 8 / 255, elems, 0, S())), 12)
rec("smooth3x3", timed(lambda: lib.advb200_smooth3x3_f32(P(x), P(out_buf), planes, H, W, S())), 8)
rec("median3x3", timed(lambda: lib.advb200_median3x3_f32(P(x), P(out_buf), planes, H, W, S())), 8)
rec("quantize16", timed(lambda: lib.advb200_quantize_f32(P(x), P(out_buf), elems, 16, S())), 8)
rec("defend_input", timed(lambda: lib.advb200_defend_input_f32(P(x), P(out_buf), planes, H, W, 16, S())), 8)
# JPEG round trip: 12 B/pixel in + 12 B/pixel out + 3 B/pixel of uint8 scratch (written, then read) = 27 B/pixel = 9 B/elem
jbytes = int(lib.advb200_jpeg_scratch_bytes(n, H, W))
jscratch = torch.empty(jbytes, dtype=torch.uint8, device=dev)
rec("jpeg_q75", timed(lambda: lib.advb200_jpeg_roundtrip_f32(P(x), P(out_buf), n, H, W, 75, P(jscratch), jbytes, S())), 9)

# CW-L2 update kernel alone (attack.py:126-137 + Adam): reads w, m, v, x0, grad_x (20 B/elem), writes w, m, v, x_adv
# (16 B/elem) = 36 B/elem; success = 0 for every image so no best_adv copy is added (that is +4 B/elem on improving steps)
per = x[0].numel()
parts = int(lib.advb200_cw_parts(per))
cw_w, cw_m, cw_v, cw_x0, cw_xa, cw_best = (torch.empty_like(x) for _ in range(6))
cw_bl2 = torch.empty(n, device=dev)
cw_bs = torch.empty(n, dtype=torch.uint8, device=dev)
cw_succ = torch.zeros(n, dtype=torch.uint8, device=dev)
cw_pa = torch.empty(n * parts + (n + 3) // 4 + 1, device=dev)
cw_pb = torch.empty_like(cw_pa)
rec("cw_init", timed(lambda: lib.advb200_cw_init_f32(P(x), P(cw_x0), P(cw_w), P(cw_m), P(cw_v), P(cw_xa), P(cw_best), P(cw_bl2),
                                                     P(cw_bs), P(cw_pa), n, per, S())), 28)
gsmall = g * 1e-3
rec("cw_step", timed(lambda: lib.advb200_cw_step_f32(P(cw_w), P(cw_m), P(cw_v), P(cw_x0), P(gsmall), P(cw_xa), P(cw_best), P(cw_bl2),
                                                     P(cw_bs), P(cw_succ), P(cw_pa), P(cw_pb), None, n, per, 0.05, 0.0316, 0.9, 0.999,
                                                     1e-8, S())), 36)

if os.environ.get("ADVB200_WORKLOADS_KERNELS_ONLY") == "1":       # just the HBM-bound kernels (seconds)
    print(json.dumps(out))
    sys.exit(0)

model = weights.seeded_resnet50(0).to(dev)
mean, std = weights.mean_std()
nm = advb200.NativeResNet50(model, precision="bf16")
y = nm.forward(x, mean, std, want_logits=False, want_preds=True)["preds"]

# CW step kernel alone (32 B/elem): time a short attack with and without ... simpler: whole CW attack rates
cw = {}
for name, kw in (("untargeted", {}), ("targeted_805", {"targeted": True, "y_target": torch.full_like(y, 805)})):
    steps = 100
    advb200.cw_l2_attack(nm, x, y, mean=mean, std=std, c=1.0, lr=0.01, steps=3, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    res, l2 = advb200.cw_l2_attack(nm, x, y, mean=mean, std=std, c=1.0, lr=0.01, steps=steps, return_l2=True, **kw)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    ok = res.success
    cw[name] = {"steps": steps, "ms": ms, "images_per_s": n / (ms * 1e-3), "success_rate": float(ok.float().mean()),
                "mean_l2_sq_of_successes": float(l2[ok].mean()) if bool(ok.any()) else None}
out["cw_l2_batch%d" % n] = cw

# defense sweep: threshold calibrated on the first 100 images (quantile 0.95), then 2 attacks x 2 eps
scores = nm.forward(x[:100], mean, std, want_logits=False, want_score=True)["score"]
thr = defenses.calibrate_threshold_from_scores(scores, 0.95)
sweep = {"threshold": thr, "rows": []}
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record()
for attack in ("fgsm", "pgd"):
    for eps in (4 / 255, 8 / 255):
        c = defenses.evaluate_defenses_batched(nm, x, y, mean, std, attack, eps, 2 / 255, 10, 1.0, 0.0, 100, 0.01, thr)
        sweep["rows"].append({"attack": attack, "eps": eps, "counters": [int(v) for v in c.cpu()]})
e1.record()
torch.cuda.synchronize()
sweep["ms_total"] = e0.elapsed_time(e1)
sweep["images_per_s_per_config"] = 4 * n / (sweep["ms_total"] * 1e-3)
out["defense_sweep_batch%d" % n] = sweep
# transfer target: VGG-19 forward (19.63 GMAC / image), batch 128 chunks
import torchvision  # noqa: E402
from advb200.transfer import NativeVGG19  # noqa: E402
torch.manual_seed(3)
vgg = NativeVGG19(torchvision.models.vgg19(weights=None).eval().to(dev))
xv = x[:128]
ms = timed(lambda: vgg.forward(xv, mean, std), reps=5, warm=2)
out["vgg19_forward_batch128"] = {"ms": ms, "images_per_s": 128 / (ms * 1e-3), "tflops": 128 * 2 * 19.63e9 / (ms * 1e-3) / 1e12}
# transformer transfer targets (forward only): ViT-B/16 (17.56 GFLOP/image nominal: 2 x 8.78 GMAC) and Swin-T (2 x 4.49 GMAC)
from advb200.transformers import NativeSwinT, NativeViTB16  # noqa: E402
for name, ctor, cls, gmac in (("vit_b_16", torchvision.models.vit_b_16, NativeViTB16, 17.56 / 2),
                              ("swin_t", torchvision.models.swin_t, NativeSwinT, 4.49)):
    torch.manual_seed(4)
    tm = cls(ctor(weights=None).eval().to(dev))
    ms = timed(lambda: tm.forward(xv, mean, std), reps=5, warm=2)
    # share of the attention kernel: time it alone through the same buffers
    out[name + "_forward_batch128"] = {"ms": ms, "images_per_s": 128 / (ms * 1e-3), "tflops": 128 * 2 * gmac * 1e9 / (ms * 1e-3) / 1e12}
print(json.dumps(out, indent=1))
