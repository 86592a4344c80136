"""Probe (build container only; imports /root/reference): the reference's OWN run-to-run spread of the CW-L2 result.
cw_l2_attack for 20 steps on 2 images of each fixture model with 8 threads, 1 thread and in float64.  Measured here:
    cond : l2 8 thr vs 1 thr 2.5e-3 / 2.8e-4, fp32 vs fp64 2.0e-3 / 5.1e-4
    seed0: l2 8 thr vs 1 thr 4.9e-4 / 1.2e-3, fp32 vs fp64 4.0e-3 / 1.9e-3
Adam's update m/sqrt(v) has unit magnitude whatever the gradient's, so a pixel whose gradient changes sign under fp32
summation noise moves by 2*lr the other way: the per-image L2 is reproducible to ~3e-3, not to the 1e-4 SURVEY section 8(d)
asks for.  tests/test_gpu_cond.py / test_gpu_configs.py assert against this floor."""
import sys, copy
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/reference")
import torch, numpy as np
from oracle import weights
import attack as ref
mean, std = weights.mean_std()
for tag in ("cond", "seed0"):
    m = weights.MODELS[tag](0)
    x = (weights.mixed_images(8, seed=77) if tag == "cond" else weights.synthetic_images(4))[:2]
    with torch.no_grad(): y = m(ref.normalize_batch(x, mean, std)).argmax(1)
    l2 = lambda a, x0: (a - x0.clamp(0,1)).flatten(1).pow(2).sum(1)
    outs = {}
    for th in (8, 1):
        torch.set_num_threads(th)
        r = ref.cw_l2_attack(m, x, y, mean=mean, std=std, c=1.0, kappa=0.0, steps=20, lr=0.01)
        outs[th] = l2(r.x_adv, x)
    torch.set_num_threads(8)
    m64 = copy.deepcopy(m).double()
    r = ref.cw_l2_attack(m64, x.double(), y, mean=mean.double(), std=std.double(), c=1.0, kappa=0.0, steps=20, lr=0.01)
    o64 = l2(r.x_adv, x.double())
    print(tag, "l2 8thr", outs[8].tolist(), "1thr", outs[1].tolist(), "fp64", o64.tolist())
    print(tag, "rel 8 vs 1:", ((outs[8]-outs[1]).abs()/outs[1]).tolist(), " fp32(8) vs fp64:", ((outs[8].double()-o64).abs()/o64).tolist(), flush=True)
