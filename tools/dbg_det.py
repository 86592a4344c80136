import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "image-recognition-adversarial-example-attack_b200")); sys.path.insert(0, str(ROOT))
import torch, advb200
from oracle import weights
dev = torch.device("cuda:0")
m = weights.seeded_resnet50(0).to(dev)
mean, std = weights.mean_std()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
x = weights.synthetic_images(n).to(dev)
y = torch.zeros(n, dtype=torch.int64, device=dev)
nm = advb200.NativeResNet50(m, precision="bf16")
outs = [nm.input_grad(x, mean, std, y, want_logits=True) for _ in range(4)]
g = [o["grad"].cpu() for o in outs]; l = [o["logits"].cpu() for o in outs]
print("logits equal:", [bool(torch.equal(l[0], t)) for t in l[1:]], "grad ndiff:", [int((g[0] != t).sum()) for t in g[1:]])
import collections
d = (g[0] != g[1]).nonzero()
print("diff by s2d col mod 13:", sorted(collections.Counter(((d[:, 3] // 2) % 13).tolist()).items()))
print("diff by s2d row mod 8:", sorted(collections.Counter(((d[:, 2] // 2) % 8).tolist()).items()))
print("diff by chan:", collections.Counter(d[:, 1].tolist()), "ph:", collections.Counter((d[:, 2] % 2).tolist()), "pw:", collections.Counter((d[:, 3] % 2).tolist()))
rd = ((g[0] - g[1]).abs() / (g[0].abs() + 1e-12))[g[0] != g[1]]
print("relative diff median", float(rd.median()), "max", float(rd.max()))
