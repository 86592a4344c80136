"""Experiment: one 256-image input-gradient pass as ONE plan on one stream vs TWO 128-image plans on two streams (the
fill / drain / tail of one half can overlap the other half's steady state).  Prints images/s for both arrangements.
Usage: python tools/two_stream_probe.py [batch] [reps]"""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "image-recognition-adversarial-example-attack_b200")); sys.path.insert(0, str(ROOT))
import torch
import advb200
from oracle import weights

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
dev = torch.device("cuda:0")
model = weights.seeded_resnet50(0).to(dev)
mean, std = (t.to(dev) for t in weights.mean_std())
x = weights.synthetic_images(64).to(dev).repeat((n + 63) // 64, 1, 1, 1)[:n].contiguous()
y = torch.zeros(n, dtype=torch.int64, device=dev)


def bench(parts):
    k = n // parts
    nms = [advb200.NativeResNet50(model, precision="bf16") for _ in range(parts)]
    streams = [torch.cuda.Stream() for _ in range(parts)]
    xs = [x[i * k:(i + 1) * k].contiguous() for i in range(parts)]
    ys = [y[i * k:(i + 1) * k].contiguous() for i in range(parts)]
    gs = [torch.empty_like(t) for t in xs]

    def once():
        for i in range(parts):
            with torch.cuda.stream(streams[i]):
                nms[i].input_grad(xs[i], mean, std, ys[i], grad_out=gs[i], inv_n_global=1.0 / n)
    for _ in range(3):
        once()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in streams:
        s.wait_event(e0)
    for _ in range(reps):
        once()
    for s in streams:
        torch.cuda.current_stream().wait_stream(s)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    return n / (ms * 1e-3), ms


for parts in (1, 2, 4):
    ips, ms = bench(parts)
    print(f"{parts} plan(s) x {n // parts} images on {parts} stream(s): {ms:.3f} ms per pass, {ips:.0f} images/s (fwd + input gradient)")
