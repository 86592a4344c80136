"""Stage-by-stage GPU bring-up probe (development aid; the real parity tests live in tests/).
Usage: python tools/gpu_probe.py <stage> ...   stages: elem conv32 convtc model32 modeltc time"""
import sys, time, os
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "image-recognition-adversarial-example-attack_b200"))
import torch
import torch.nn.functional as F
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
import advb200
from advb200 import _lib
from ctypes import c_void_p

dev = torch.device("cuda:0")


def rel(a, b):
    a = a.double(); b = b.double()
    return float((a - b).norm() / (b.norm() + 1e-30))


def stage_elem():
    g = torch.Generator(device="cpu").manual_seed(0)
    x0 = torch.rand(4, 3, 224, 224, generator=g)
    x = (x0 + torch.empty_like(x0).uniform_(-8 / 255, 8 / 255, generator=g)).clamp(0, 1)
    gr = torch.randn(4, 3, 224, 224, generator=g)
    gr[0, 0, 0, :5] = 0
    eps, alpha = 8 / 255, 2 / 255
    ref = x + alpha * gr.sign(); ref = torch.max(torch.min(ref, x0 + eps), x0 - eps).clamp(0, 1)
    xa = x.to(dev).clone()
    advb200.attack._linf_step(xa, x0.to(dev), gr.to(dev), alpha, eps, True)
    print("pgd step bit-exact:", torch.equal(xa.cpu(), ref))
    ref2 = (x + eps * gr.sign()).clamp(0, 1)
    xa = x.to(dev).clone()
    advb200.attack._linf_step(xa, None, gr.to(dev), eps, eps, False)
    print("fgsm step bit-exact:", torch.equal(xa.cpu(), ref2))
    y = torch.rand(4, 3, 224, 224, generator=g) * 1.2 - 0.1
    q = advb200._defense_quantization(y.to(dev)).cpu()
    qr = torch.round(y.clamp(0, 1) * 15) / 15
    print("quantize bit-exact:", torch.equal(q, qr), float((q - qr).abs().max()))
    s = advb200._defense_smoothing(y.to(dev)).cpu()
    sr = F.avg_pool2d(y, 3, 1, 1)
    print("smooth bit-exact:", torch.equal(s, sr), float((s - sr).abs().max()))
    d = advb200.defend_input(y.to(dev)).cpu()
    t = F.avg_pool2d(y.clamp(0, 1), 3, 1, 1)
    dr = (torch.round(t.clamp(0, 1) * 15) / 15).clamp(0, 1)
    print("defend bit-exact:", torch.equal(d, dr), float((d - dr).abs().max()))


SHAPES = [  # n,h,w,cin,cout,k,stride
    (2, 56, 56, 64, 64, 1, 1), (2, 56, 56, 64, 256, 1, 1), (2, 56, 56, 64, 64, 3, 1),
    (2, 56, 56, 128, 128, 3, 2), (2, 56, 56, 256, 512, 1, 2), (3, 28, 28, 128, 128, 3, 1),
    (5, 14, 14, 256, 256, 3, 1), (5, 14, 14, 512, 512, 3, 2), (5, 7, 7, 512, 2048, 1, 1),
    (3, 7, 7, 512, 512, 3, 1), (2, 14, 14, 1024, 2048, 1, 2),
]


def run_conv(prec, x_nhwc, w, bias, add, mask, n, h, wd, cin, cout, k, stride, relu, dgrad, out_shape):
    lib = _lib.load()
    dt = torch.bfloat16 if prec == 0 else torch.float32
    out = torch.full(out_shape, float("nan"), dtype=dt, device=dev)
    rc = lib.advb200_test_conv(prec, _lib.ptr(x_nhwc), _lib.ptr(w), _lib.ptr(bias), _lib.ptr(add), _lib.ptr(mask),
                               _lib.ptr(out), n, h, wd, cin, cout, k, stride, relu, dgrad, _lib.stream_ptr())
    _lib.check(rc, "test_conv")
    torch.cuda.synchronize()
    return out


def stage_conv(prec):
    dt = torch.bfloat16 if prec == 0 else torch.float32
    g = torch.Generator(device="cpu").manual_seed(1)
    for (n, h, wd, cin, cout, k, stride) in SHAPES:
        ho, wo = h // stride, wd // stride
        x = torch.randn(n, cin, h, wd, generator=g).to(dev)
        w = (torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).to(dev)
        b = torch.randn(cout, generator=g).to(dev)
        res = torch.randn(n, cout, ho, wo, generator=g).to(dev)
        xq = x.to(dt).float(); wq = w.to(dt).float() if prec == 0 else w; rq = res.to(dt).float()
        ref = F.relu(F.conv2d(xq, wq, b, stride=stride, padding=k // 2) + rq)
        out = run_conv(prec, xq.permute(0, 2, 3, 1).contiguous().to(dt), w, b,
                       rq.permute(0, 2, 3, 1).contiguous().to(dt), None, n, h, wd, cin, cout, k, stride, 1, 0,
                       (n, ho, wo, cout))
        e1 = rel(out.float().permute(0, 3, 1, 2), ref)
        # dgrad
        gy = torch.randn(n, cout, ho, wo, generator=g).to(dev)
        gyq = gy.to(dt).float()
        act = torch.randn(n, cin, h, wd, generator=g).to(dev)
        addg = torch.randn(n, cin, h, wd, generator=g).to(dev).to(dt).float()
        refg = torch.nn.grad.conv2d_input((n, cin, h, wd), wq, gyq, stride=stride, padding=k // 2)
        plain = (k == 1 and stride == 2)   # downsample dgrad: the engine uses it without add / mask
        if not plain:
            refg = (refg + addg) * (act.to(dt).float() > 0)
        outg = run_conv(prec, gyq.permute(0, 2, 3, 1).contiguous().to(dt), w, None,
                        None if plain else addg.permute(0, 2, 3, 1).contiguous().to(dt),
                        None if plain else act.permute(0, 2, 3, 1).contiguous().to(dt), n, h, wd, cin, cout, k,
                        stride, 0, 1, (n, h, wd, cin))
        e2 = rel(outg.float().permute(0, 3, 1, 2), refg)
        print(f"prec={prec} conv n{n} {h}x{wd} {cin}->{cout} k{k} s{stride}: fwd rel {e1:.2e}  dgrad rel {e2:.2e}"
              f"  nan fwd {int(torch.isnan(out.float()).sum())} dgrad {int(torch.isnan(outg.float()).sum())}",
              flush=True)


def make_model():
    import torchvision
    torch.manual_seed(0)
    return torchvision.models.resnet50(weights=None).eval()


def stage_model(precision, n=2):
    m = make_model().to(dev)
    g = torch.Generator().manual_seed(1234)
    x = torch.rand(n, 3, 224, 224, generator=g).to(dev)
    mean = torch.tensor([0.485, 0.456, 0.406]); std = torch.tensor([0.229, 0.224, 0.225])
    nm = advb200.NativeResNet50(m, precision=precision)
    xr = x.clone().requires_grad_(True)
    logits = m(advb200.normalize_batch(xr, mean, std))
    y = logits.argmax(1)
    loss = F.cross_entropy(logits, y)
    gref = torch.autograd.grad(loss, xr)[0]
    out = nm.input_grad(x, mean, std, y, loss="ce", want_logits=True, want_preds=True)
    torch.cuda.synchronize()
    print(precision, "logits rel", rel(out["logits"], logits.detach()), "preds eq", torch.equal(out["preds"], y),
          "pred", y.tolist())
    print(precision, "grad rel", rel(out["grad"], gref), "max|g| ref", float(gref.abs().max()), "ours",
          float(out["grad"].abs().max()), "nan", int(torch.isnan(out["grad"]).sum()))
    mask = gref.abs() > 1e-6 * gref.abs().max()
    agree = (out["grad"].sign() == gref.sign())[mask].float().mean()
    print(precision, "sign agreement on mask", float(agree))
    # fp64 truth
    m64 = make_model().double().to(dev)
    x64 = x.double().clone().requires_grad_(True)
    l64 = m64(advb200.normalize_batch(x64, mean.double(), std.double()))
    g64 = torch.autograd.grad(F.cross_entropy(l64, y), x64)[0]
    print(precision, "vs fp64: ours grad rel", rel(out["grad"], g64), " torch-fp32 grad rel", rel(gref, g64),
          " ours logits rel", rel(out["logits"], l64.detach()), " torch logits rel", rel(logits.detach(), l64.detach()))
    sc = nm.forward(x, mean, std, want_logits=False, want_score=True, stop_at_layer3=True)["score"]
    f = m.layer3(m.layer2(m.layer1(m.maxpool(m.relu(m.bn1(m.conv1(advb200.normalize_batch(x, mean, std))))))))
    sref = (f.pow(2).sum(dim=[1, 2, 3]).sqrt() / 1024 + f.var(dim=[2, 3]).mean(dim=1) * 0.1).clamp(0, 100)
    print(precision, "detector score ours", sc.tolist(), "ref", sref.tolist())


def stage_time(precision="bf16", n=256, steps=10, warm=2):
    m = make_model().to(dev)
    g = torch.Generator().manual_seed(1234)
    x = torch.rand(n, 3, 224, 224, generator=g).to(dev)
    mean = torch.tensor([0.485, 0.456, 0.406]); std = torch.tensor([0.229, 0.224, 0.225])
    nm = advb200.NativeResNet50(m, precision=precision)
    y = nm.forward(x, mean, std, want_logits=False, want_preds=True)["preds"]
    for _ in range(warm):
        advb200.pgd_linf_attack(nm, x, y, eps=8 / 255, alpha=2 / 255, steps=1, mean=mean, std=std)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    xa = advb200.pgd_linf_attack(nm, x, y, eps=8 / 255, alpha=2 / 255, steps=steps, mean=mean, std=std)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    pa = nm.forward(xa, mean, std, want_logits=False, want_preds=True)["preds"]
    print(f"{precision} PGD-{steps} N={n}: {ms:.1f} ms -> {n / ms * 1e3:.1f} img/s; success {(pa != y).float().mean():.3f}")


if __name__ == "__main__":
    for st in sys.argv[1:]:
        print("=== stage", st, flush=True)
        t0 = time.time()
        {"elem": stage_elem, "conv32": lambda: stage_conv(1), "convtc": lambda: stage_conv(0),
         "model32": lambda: stage_model("fp32"), "modeltc": lambda: stage_model("bf16"),
         "time": stage_time, "time2": lambda: stage_time("bf16", 256, 2, 1), "time32": lambda: stage_time("fp32", 32, 2)}[st]()
        print(f"=== stage {st} done in {time.time() - t0:.1f}s", flush=True)
