"""Drop-in for the reference's blackbox_transfer.py on the native kernels: adversarials crafted on ResNet-50
(FGSM / PGD / CW-L2, blackbox_transfer.py:172-214) are classified by the transfer targets and the attack-success
table of :316-340 is printed in the same format.  Batched (the reference loops image by image); the success
criterion is the reference's: the target's prediction on the adversarial differs from its own on the clean image.

Targets: VGG-19 (advb200.transfer.NativeVGG19), ViT-B/16 and Swin-T (advb200.transformers) all run natively on the
device.  Plotting (--visualize_n) is out of scope.

    python blackbox_transfer.py --image_dir picture --attacks fgsm pgd cw
    python blackbox_transfer.py --synthetic 64 --weights random --attacks fgsm pgd       # offline smoke run
"""
from __future__ import annotations

import argparse
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402
import torchvision.models as tv_models  # noqa: E402
from PIL import Image  # noqa: E402

from advb200 import cw_l2_attack, fgsm_attack, native_model, pgd_linf_attack, preprocess, set_default_precision  # noqa: E402
from advb200.transfer import NativeVGG19, transfer_counts  # noqa: E402
from advb200.transformers import NativeSwinT, NativeViTB16  # noqa: E402


def build_parser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser()
    parser.add_argument("--image_dir", type=str, default="picture")
    parser.add_argument("--attacks", type=str, nargs="+", default=["fgsm", "pgd", "cw"], choices=["fgsm", "pgd", "cw"])
    parser.add_argument("--eps", type=float, default=8 / 255)
    parser.add_argument("--alpha", type=float, default=2 / 255)
    parser.add_argument("--steps", type=int, default=10)
    parser.add_argument("--cw_c", type=float, default=1.0)
    parser.add_argument("--cw_kappa", type=float, default=0.0)
    parser.add_argument("--cw_steps", type=int, default=200)
    parser.add_argument("--cw_lr", type=float, default=0.01)
    parser.add_argument("--visualize_n", type=int, default=3)
    # extras
    parser.add_argument("--weights", choices=["pretrained", "random"], default="pretrained")
    parser.add_argument("--seed", type=int, default=0)
    parser.add_argument("--synthetic", type=int, default=0, help="use N seeded synthetic images instead of --image_dir")
    parser.add_argument("--batch", type=int, default=128)
    parser.add_argument("--precision", choices=["bf16", "fp32"], default="bf16")
    return parser


def load_models(kind: str, seed: int, device):
    if kind == "random":
        torch.manual_seed(seed)
        src = tv_models.resnet50(weights=None).eval()
        torch.manual_seed(seed + 1)
        vgg = tv_models.vgg19(weights=None).eval()
    else:
        try:
            src = tv_models.resnet50(weights=tv_models.ResNet50_Weights.DEFAULT).eval()
            vgg = tv_models.vgg19(weights=tv_models.VGG19_Weights.IMAGENET1K_V1).eval()
        except Exception as e:  # offline box
            raise SystemExit(f"could not load the pretrained torchvision checkpoints ({e}); use --weights random")
    return src.to(device), vgg.to(device)


def load_transformers(kind: str, seed: int, device):
    """blackbox_transfer.py:36-54 `load_vit` / `load_swin` (IMAGENET1K_V1), or seeded random init offline.  torchvision
    zero-initialises the ViT classifier head; the random-init variant redraws it so the column is not degenerate."""
    if kind == "random":
        torch.manual_seed(seed + 2)
        vit = tv_models.vit_b_16(weights=None).eval()
        with torch.no_grad():
            vit.heads.head.weight.normal_(std=0.02)
        torch.manual_seed(seed + 3)
        swin = tv_models.swin_t(weights=None).eval()
    else:
        try:
            vit = tv_models.vit_b_16(weights=tv_models.ViT_B_16_Weights.IMAGENET1K_V1).eval()
            swin = tv_models.swin_t(weights=tv_models.Swin_T_Weights.IMAGENET1K_V1).eval()
        except Exception as e:  # offline box
            raise SystemExit(f"could not load the pretrained torchvision checkpoints ({e}); use --weights random")
    return vit.to(device), swin.to(device)


def iter_images(image_dir: Path):
    exts = (".jpg", ".jpeg", ".png", ".bmp")
    for path in sorted(image_dir.iterdir()):
        if path.is_file() and path.suffix.lower() in exts:
            yield path


def main(argv=None) -> None:
    args = build_parser().parse_args(argv)
    if not torch.cuda.is_available():
        raise SystemExit("this CLI runs on a CUDA device (B200); no CPU fallback")
    device = torch.device("cuda")
    set_default_precision(args.precision)
    src, vgg = load_models(args.weights, args.seed, device)
    vit, swin = load_transformers(args.weights, args.seed, device)
    source = native_model(src)
    targets = {"VGG19": NativeVGG19(vgg), "ViT": NativeViTB16(vit), "Swin": NativeSwinT(swin)}
    mean = torch.tensor([0.485, 0.456, 0.406], device=device)
    std = torch.tensor([0.229, 0.224, 0.225], device=device)

    if args.synthetic > 0:
        g = torch.Generator().manual_seed(1234)
        images = torch.rand(args.synthetic, 3, 224, 224, generator=g).to(device)
    else:
        image_dir = Path(args.image_dir)
        if not image_dir.is_dir():
            raise SystemExit(f"image_dir not found: {image_dir}")
        tf = preprocess.get_transform(device)
        xs = [tf(Image.open(p).convert("RGB")) for p in iter_images(image_dir)]
        if not xs:
            raise SystemExit(f"No image files found in directory: {image_dir}")
        images = torch.stack(xs)

    stats = {a: {t: {"success": 0, "total": 0} for t in targets} for a in ("fgsm", "pgd", "cw")}
    for lo in range(0, images.shape[0], int(args.batch)):
        x = images[lo:lo + int(args.batch)]
        y_source = source.forward(x, mean, std, want_logits=False, want_preds=True)["preds"]
        for name in args.attacks:
            if name == "fgsm":
                x_adv = fgsm_attack(source, x, y_source, eps=float(args.eps), mean=mean, std=std)
            elif name == "pgd":
                x_adv = pgd_linf_attack(source, x, y_source, eps=float(args.eps), alpha=float(args.alpha),
                                        steps=int(args.steps), mean=mean, std=std)
            else:
                x_adv = cw_l2_attack(source, x, y_source, mean=mean, std=std, c=float(args.cw_c),
                                     kappa=float(args.cw_kappa), steps=int(args.cw_steps), lr=float(args.cw_lr),
                                     targeted=False, y_target=None).x_adv
            for tname, target in targets.items():
                s, t = transfer_counts(target, x, x_adv, mean, std)
                stats[name][tname]["success"] += s
                stats[name][tname]["total"] += t

    # blackbox_transfer.py:316-340 — rows = attacks, columns = target models
    print("\t".join(["Attack/Model", "VGG19", "ViT", "Swin"]))
    for name in ("fgsm", "pgd", "cw"):
        row = [name.upper()]
        for tname in ("VGG19", "ViT", "Swin"):
            total, success = stats[name][tname]["total"], stats[name][tname]["success"]
            row.append(f"{0.0 if total == 0 else 100.0 * success / total:.1f}%")
        print("\t".join(row))


if __name__ == "__main__":
    main()
