#!/usr/bin/env python
"""Drop-in for the reference's ResNet.py CLI (same flags, same stdout format) with the attack loop
on the B200 through libadvb200.

    python ResNet.py <image|dir> --attack fgsm|pgd|cw --topk 5 [--eps --alpha --steps --cw_* --target --save_adv]

Extra flags (not in the reference): --weights pretrained|random, --seed, --precision bf16|fp32.
The reference always loads torchvision's pretrained checkpoint (ResNet.py:13), which needs network
access; `--weights random --seed 0` builds the fixed random-init model the parity tests use.
"""
from __future__ import annotations

import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch  # noqa: E402
import torchvision.models as models  # noqa: E402
import torchvision.transforms as transforms  # noqa: E402
from PIL import Image  # noqa: E402

from advb200 import cw_l2_attack, fgsm_attack, native_model, pgd_linf_attack, preprocess, set_default_precision  # noqa: E402


def build_parser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser()
    parser.add_argument("image", nargs="?", default="example.jpg")
    parser.add_argument("--topk", type=int, default=5)
    parser.add_argument("--attack", choices=["none", "fgsm", "pgd", "cw"], default="none")
    parser.add_argument("--label", type=int, default=None)
    parser.add_argument("--eps", type=float, default=8 / 255)
    parser.add_argument("--alpha", type=float, default=2 / 255)
    parser.add_argument("--steps", type=int, default=10)
    parser.add_argument("--cw_c", type=float, default=1.0)
    parser.add_argument("--cw_kappa", type=float, default=0.0)
    parser.add_argument("--cw_steps", type=int, default=1000)
    parser.add_argument("--cw_lr", type=float, default=0.01)
    parser.add_argument("--target", type=int, default=None)
    parser.add_argument("--save_adv", type=str, default=None)
    # extras
    parser.add_argument("--weights", choices=["pretrained", "random"], default="pretrained")
    parser.add_argument("--seed", type=int, default=0)
    parser.add_argument("--precision", choices=["bf16", "fp32"], default="bf16")
    return parser


def load_model(kind: str, seed: int):
    if kind == "random":
        torch.manual_seed(seed)
        return models.resnet50(weights=None).eval()
    try:
        return models.resnet50(weights=models.ResNet50_Weights.IMAGENET1K_V1).eval()   # == pretrained=True
    except Exception as e:  # offline box
        raise SystemExit(f"could not load the pretrained torchvision checkpoint ({e}); use --weights random")


def main(argv=None) -> None:
    args = build_parser().parse_args(argv)
    if not torch.cuda.is_available():
        raise SystemExit("this CLI runs the attack path on a CUDA device (B200); no CPU fallback")
    device = torch.device("cuda")
    set_default_precision(args.precision)
    model = load_model(args.weights, args.seed).to(device)
    nm = native_model(model)
    mean = torch.tensor([0.485, 0.456, 0.406])
    std = torch.tensor([0.229, 0.224, 0.225])
    transform = preprocess.get_transform(device)      # Resize(256) -> CenterCrop(224) -> ToTensor, on the device
    categories = None
    try:
        categories = models.ResNet50_Weights.DEFAULT.meta.get("categories")
    except Exception:
        categories = None

    def report(logits, topk):
        prob = torch.softmax(logits, dim=1)
        top_prob, top_idx = prob.topk(topk, dim=1)
        for rank, (p, idx) in enumerate(zip(top_prob[0].cpu().tolist(), top_idx[0].cpu().tolist()), start=1):
            label = categories[idx] if categories and idx < len(categories) else str(idx)
            print(f"Top {rank}: {label} (class {idx}), prob = {p:.4f}")

    def run_for_image(image_path: Path):
        if not image_path.is_file():
            print(f"Image file not found: {image_path}")
            return
        x = transform(Image.open(image_path).convert("RGB")).unsqueeze(0)
        logits = nm.forward(x, mean, std)["logits"]
        pred = logits.argmax(dim=1)
        topk = max(1, int(args.topk))
        print(f"Image: {image_path}")
        print("Clean:")
        report(logits, topk)
        if args.attack == "none":
            return
        y_true = torch.tensor([args.label], device=device) if args.label is not None else pred.detach()
        if args.attack == "fgsm":
            x_adv = fgsm_attack(nm, x, y_true, eps=float(args.eps), mean=mean, std=std)
        elif args.attack == "pgd":
            x_adv = pgd_linf_attack(nm, x, y_true, eps=float(args.eps), alpha=float(args.alpha),
                                    steps=int(args.steps), mean=mean, std=std)
        else:
            targeted = args.target is not None
            y_target = torch.tensor([int(args.target)], device=device) if targeted else None
            x_adv = cw_l2_attack(nm, x, y_true, mean=mean, std=std, c=float(args.cw_c), kappa=float(args.cw_kappa),
                                 steps=int(args.cw_steps), lr=float(args.cw_lr), targeted=targeted,
                                 y_target=y_target).x_adv
        print(f"Adversarial ({args.attack}):")
        report(nm.forward(x_adv, mean, std)["logits"], topk)
        if args.save_adv:
            out_path = Path(args.save_adv)
            if out_path.parent:
                out_path.parent.mkdir(parents=True, exist_ok=True)
            preprocess.to_pil_image(x_adv[0].detach()).save(out_path)

    image_path = Path(args.image)
    if image_path.is_dir():
        image_files = []
        for pattern in ("*.jpg", "*.jpeg", "*.png", "*.bmp"):
            image_files.extend(sorted(image_path.glob(pattern)))
        if not image_files:
            raise SystemExit(f"No image files found in directory: {image_path}")
        for img_path in image_files:
            run_for_image(img_path)
    else:
        if not image_path.is_file():
            raise SystemExit(
                f"Image file not found: {image_path}. "
                f"Place an image at '{Path('example.jpg').resolve()}' or pass a path: python ResNet.py <image_path>")
        run_for_image(image_path)


if __name__ == "__main__":
    main()
