"""One warm pass at batch N inside a cudaProfilerStart/Stop range (ncu --profile-from-start off captures only it,
not the plan-time autotuner's launches):
    python tools/one_pass.py [N] [mode]     mode = grad (forward + input gradient, default) | pgd (a 2-step PGD attack)"""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "image-recognition-adversarial-example-attack_b200")); sys.path.insert(0, str(ROOT))
import torch
import advb200
from oracle import weights
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
mode = sys.argv[2] if len(sys.argv) > 2 else "grad"
dev = torch.device("cuda:0")
nm = advb200.NativeResNet50(weights.seeded_resnet50(0).to(dev), precision="bf16")
nm.max_chunk = n
mean, std = weights.mean_std()
x = weights.synthetic_images(min(n, 32)).to(dev)
x = x.repeat((n + x.shape[0] - 1) // x.shape[0], 1, 1, 1)[:n].contiguous()
y = torch.zeros(n, dtype=torch.int64, device=dev)
g = torch.empty_like(x)


def one():
    if mode == "pgd":
        advb200.pgd_linf_attack(nm, x, y, eps=8 / 255, alpha=2 / 255, steps=2, mean=mean, std=std)
    else:
        nm.input_grad(x, mean, std, y, grad_out=g)


for _ in range(2):
    one()
torch.cuda.synchronize()
torch.cuda.profiler.start()
one()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
