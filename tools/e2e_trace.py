"""Debug: per-step device timeline of the host-buffer pipeline (where does an occasional slow end-to-end run lose
its time?).  python tools/e2e_trace.py [runs]"""
import functools, itertools, sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "image-recognition-adversarial-example-attack_b200")); sys.path.insert(0, str(ROOT))
import torch
import advb200
from oracle import weights
runs = int(sys.argv[1]) if len(sys.argv) > 1 else 8
dev = torch.device("cuda:0")
nm = advb200.NativeResNet50(weights.seeded_resnet50(0).to(dev), precision="bf16")
mean, std = weights.mean_std()
n = 256
x = weights.synthetic_images(32).to(dev).repeat(8, 1, 1, 1).contiguous()
y = nm.forward(x, mean, std, want_logits=False, want_preds=True)["preds"]
marks = []
def attack(xd):
    e0 = torch.cuda.Event(enable_timing=True); e0.record()
    t0 = time.perf_counter()
    r = advb200.pgd_linf_attack(nm, xd, y, eps=8 / 255, alpha=2 / 255, steps=10, mean=mean, std=std)
    t1 = time.perf_counter()
    e1 = torch.cuda.Event(enable_timing=True); e1.record()
    marks.append((e0, e1, t0, t1))
    return r
for _ in range(3):
    attack(x)
torch.cuda.synchronize()
xh = x.cpu().pin_memory()
outs = [torch.empty_like(xh).pin_memory() for _ in range(2)]
pipe = advb200.HostBatchPipeline(attack, dev)
for _ in pipe.run([xh, xh], itertools.cycle(outs)):
    pass
for r in range(runs):
    torch.cuda.synchronize()
    marks.clear()
    s0 = torch.cuda.Event(enable_timing=True); s1 = torch.cuda.Event(enable_timing=True)
    w0 = time.perf_counter()
    s0.record()
    for _ in pipe.run((xh for _ in range(6)), itertools.cycle(outs)):
        pass
    torch.cuda.current_stream().wait_stream(pipe.d2h)
    s1.record()
    torch.cuda.synchronize()
    tot = s0.elapsed_time(s1)
    starts = [s0.elapsed_time(m[0]) for m in marks]
    durs = [m[0].elapsed_time(m[1]) for m in marks]
    host = [(m[2] - w0) * 1e3 for m in marks]
    hdur = [(m[3] - m[2]) * 1e3 for m in marks]
    print(f"run {r}: total {tot:.1f} ms  ({n * 6 / tot * 1e3:.0f} img/s)")
    print("   attack start (dev ms):", [round(v, 1) for v in starts])
    print("   attack dur   (dev ms):", [round(v, 1) for v in durs])
    print("   host enqueue at  (ms):", [round(v, 1) for v in host], " host enqueue took:", [round(v, 1) for v in hdur])
