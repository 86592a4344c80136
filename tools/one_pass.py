"""One warm forward + input-gradient pass at batch N (for ncu captures): python tools/one_pass.py [N] [passes]"""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "image-recognition-adversarial-example-attack_b200")); sys.path.insert(0, str(ROOT))
import torch
import advb200
from oracle import weights
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
passes = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dev = torch.device("cuda:0")
nm = advb200.NativeResNet50(weights.seeded_resnet50(0).to(dev), precision="bf16")
nm.max_chunk = n
mean, std = weights.mean_std()
x = weights.synthetic_images(min(n, 32)).to(dev)
x = x.repeat((n + x.shape[0] - 1) // x.shape[0], 1, 1, 1)[:n].contiguous()
y = torch.zeros(n, dtype=torch.int64, device=dev)
g = torch.empty_like(x)
for _ in range(passes):
    nm.input_grad(x, mean, std, y, grad_out=g)
torch.cuda.synchronize()
print("ok")
