"""One pass over every kernel family of libadvb200.so with little else on the device, for compute-sanitizer
(tools/gpu_sanitize.sh): the whole-network forward + input gradient in both precisions (all conv launch kinds of the
real plan incl. the fused bottleneck tails and the multi-class stride-2 launches), the three attacks, the defenses, the
JPEG round trip, the preprocessing kernels, the batched evaluator and the VGG-19 transfer forward."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "image-recognition-adversarial-example-attack_b200"))
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402
import advb200  # noqa: E402
from advb200 import _lib  # noqa: E402
from oracle import weights  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
dev = torch.device("cuda:0")
model = weights.conditioned_resnet50(0).to(dev)
mean, std = weights.mean_std()
x = weights.mixed_images(n, seed=5).to(dev)
before = _lib.launch_count()
for precision in ("bf16", "fp32"):
    nm = advb200.NativeResNet50(model, precision=precision)
    y = nm.forward(x, mean, std, want_preds=True)["preds"]
    nm.input_grad(x, mean, std, y, loss="ce", want_logits=True)
    advb200.fgsm_attack(nm, x, y, eps=8 / 255, mean=mean, std=std)
    advb200.pgd_linf_attack(nm, x, y, eps=8 / 255, alpha=2 / 255, steps=2, mean=mean, std=std)
    advb200.cw_l2_attack(nm, x, y, mean=mean, std=std, steps=2, lr=0.01)
    advb200.cw_l2_attack(nm, x, y, mean=mean, std=std, steps=2, lr=0.01, targeted=True, y_target=torch.full((n,), 805, device=dev))
    advb200.evaluate_defenses_batched(nm, x, y, mean, std, "fgsm", 4 / 255, 2 / 255, 2, 1.0, 0.0, 2, 0.01, 0.4)
    if precision == "bf16":
        xa = x.clone()
        nm.attack_step(xa, x, mean, std, y, alpha=2 / 255, eps=8 / 255)
    torch.cuda.synchronize()
for shape in ((n, 3, 224, 224), (2, 3, 17, 23), (1, 1, 1, 1)):
    t = torch.rand(shape, device=dev)
    advb200.defend_input(t); advb200._defense_smoothing(t); advb200._defense_quantization(t, 16)
    advb200._defense_median_smoothing(t)
advb200._jpeg_compress_batch(torch.rand(2, 3, 37, 51, device=dev), 75)
advb200._jpeg_compress_batch(x[:1], 30)
from advb200 import preprocess  # noqa: E402
img = (torch.rand(300, 517, 3) * 255).to(torch.uint8).numpy()
from PIL import Image  # noqa: E402
t = preprocess.get_transform(dev)(Image.fromarray(img))
preprocess.to_pil_image(t)
try:
    import torchvision
    torch.manual_seed(0)
    from advb200.transfer import NativeVGG19
    vgg = NativeVGG19(torchvision.models.vgg19(weights=None).eval().to(dev))
    vgg.forward(x[:2], mean, std)
except Exception as e:  # the transfer target is optional for this pass
    print("vgg19 skipped:", e)
torch.cuda.synchronize()
print(f"sanitize pass: {_lib.launch_count() - before} native kernel launches, no error reported by the runtime")
