import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "image-recognition-adversarial-example-attack_b200")); sys.path.insert(0, str(ROOT))
import torch, advb200
from oracle import weights
dev = torch.device("cuda:0")
m = weights.seeded_resnet50(0).to(dev)
mean, std = weights.mean_std()
x = weights.synthetic_images(2).to(dev)
y = torch.zeros(2, dtype=torch.int64, device=dev)
g = {}
for prec in ("fp32", "bf16"):
    nm = advb200.NativeResNet50(m, precision=prec)
    g[prec] = nm.input_grad(x, mean, std, y)["grad"].cpu()
    g2 = nm.input_grad(x, mean, std, y)["grad"].cpu()
    print(prec, "deterministic:", bool(torch.equal(g[prec], g2)), "ndiff", int((g[prec] != g2).sum()))
a, b = g["fp32"], g["bf16"]
print("max |fp32|", float(a.abs().max()), "max |bf16|", float(b.abs().max()))
err = (a - b).abs()
bad = err > 10 * a.abs().max()
print("bad count", int(bad.sum()), "of", bad.numel())
idx = bad.nonzero()
print(idx[:20].tolist())
if len(idx):
    import collections
    print("cols", sorted(collections.Counter((idx[:, 3] // 2).tolist()).items())[:40])
    print("rows", sorted(collections.Counter((idx[:, 2] // 2).tolist()).items())[:40])
    print("chan", collections.Counter(idx[:, 1].tolist()), "parity h", collections.Counter((idx[:, 2] % 2).tolist()), "w", collections.Counter((idx[:, 3] % 2).tolist()))
rel = float((a - b).norm() / a.norm())
print("rel", rel)
