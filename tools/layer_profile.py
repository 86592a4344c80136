"""Per-launch profile of the tcgen05 conv launches of one forward + input-gradient pass.
Usage: python tools/layer_profile.py [batch] [reps] > gpurun_out/layers.csv
Every conv launch is bracketed by CUDA events on its stream (advb200_profile_enable); the table
lists algorithmic TFLOP/s and GB/s per launch and the fraction of the roofline bound
max(flops/peak_tf, bytes/peak_bw) it reaches."""
import ctypes
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "image-recognition-adversarial-example-attack_b200"))
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402
import advb200  # noqa: E402
from advb200 import _lib  # noqa: E402
from oracle import weights  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").is_file() else {}
PTF = float(peaks.get("bf16_tflops", 1590.0))
PBW = float(peaks.get("hbm_gbs", 6650.0))

dev = torch.device("cuda:0")
model = weights.seeded_resnet50(0).to(dev)
mean, std = weights.mean_std()
nm = advb200.NativeResNet50(model, precision="bf16")
nm.max_chunk = n
x = weights.synthetic_images(min(n, 64)).to(dev)
x = x.repeat((n + x.shape[0] - 1) // x.shape[0], 1, 1, 1)[:n].contiguous()
y = torch.zeros(n, dtype=torch.int64, device=dev)
g = torch.empty_like(x)
for _ in range(2):
    nm.input_grad(x, mean, std, y, grad_out=g)
torch.cuda.synchronize()
lib = _lib.load()
acc = {}
order = []
for _ in range(reps):
    lib.advb200_profile_enable(1)
    nm.input_grad(x, mean, std, y, grad_out=g)
    torch.cuda.synchronize()
    need = lib.advb200_profile_dump(None, 0)
    lib.advb200_profile_enable(1)
    nm.input_grad(x, mean, std, y, grad_out=g)
    torch.cuda.synchronize()
    buf = ctypes.create_string_buffer(int(need) + 4096)
    lib.advb200_profile_dump(buf, len(buf))
    lib.advb200_profile_enable(0)
    for line in buf.value.decode().strip().split("\n")[1:]:
        name, bn, tiles, ksteps, stages, ms, fl, by = line.split(",")
        if name not in acc:
            acc[name] = [bn, int(tiles), int(ksteps), int(stages), [], float(fl), float(by)]
            order.append(name)
        acc[name][4].append(float(ms))

print("name,bn,tiles,ksteps,stages,us,tflops,gbs,bound,frac_of_bound")
tot_us = tot_min = 0.0
for name in order:
    bn, tiles, ksteps, stages, mss, fl, by = acc[name]
    ms = sorted(mss)[len(mss) // 2]
    t_tf, t_bw = fl / (PTF * 1e12), by / (PBW * 1e9)
    tmin = max(t_tf, t_bw)
    tot_us += ms * 1e3
    tot_min += tmin * 1e6
    print(f"{name},{bn},{tiles},{ksteps},{stages},{ms*1e3:.1f},{fl/ms/1e9:.0f},{by/ms/1e6:.0f},"
          f"{'tensor' if t_tf >= t_bw else 'hbm'},{tmin/(ms*1e-3):.3f}")
print(f"# batch {n}: conv launches total {tot_us:.0f} us, roofline bound {tot_min:.0f} us "
      f"(peaks {PTF} TF/s, {PBW} GB/s)")
