#!/usr/bin/env python
"""Drop-in for the reference's defense_experiments.py sweep (same flags, same summary lines), batched
and sharded: every rank (one per GPU under torchrun) attacks a contiguous shard of the images and the
int64[7] counter vector is summed with one all-reduce per (attack, eps).

Extra flags: --synthetic N (seeded U[0,1) 224x224 images instead of files), --batch, --precision,
--weights pretrained|random, --seed.  Plotting (reference lines 350-530, 852-872) is out of scope.
"""
from __future__ import annotations

import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch  # noqa: E402
import torchvision.models as models  # noqa: E402
import torchvision.transforms as transforms  # noqa: E402

import advb200  # noqa: E402
from advb200 import defenses, dist as adist, preprocess  # noqa: E402


def build_parser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser()
    parser.add_argument("--model_type", type=str, choices=["standard", "robust"], default="standard")
    parser.add_argument("--image_dir", type=str, default=None)
    parser.add_argument("--image", type=str, default="example.jpg")
    parser.add_argument("--attacks", type=str, nargs="+", default=["fgsm", "pgd", "cw"], choices=["fgsm", "pgd", "cw"])
    parser.add_argument("--eps_list", type=float, nargs="+", default=[4 / 255, 8 / 255, 16 / 255])
    parser.add_argument("--alpha", type=float, default=2 / 255)
    parser.add_argument("--steps", type=int, default=10)
    parser.add_argument("--cw_c", type=float, default=1.0)
    parser.add_argument("--cw_kappa", type=float, default=0.0)
    parser.add_argument("--cw_steps", type=int, default=100)
    parser.add_argument("--cw_lr", type=float, default=0.01)
    parser.add_argument("--detector_threshold", type=float, default=None)
    parser.add_argument("--calibrate_dir", type=str, default=None)
    parser.add_argument("--calibrate_n", type=int, default=100)
    parser.add_argument("--calibrate_quantile", type=float, default=0.95)
    parser.add_argument("--use_jpeg", action="store_true")
    parser.add_argument("--jpeg_quality", type=int, default=75)
    parser.add_argument("--output_dir", type=str, default="./defense_results")
    parser.add_argument("--viz_samples", type=int, default=5)
    # extras
    parser.add_argument("--synthetic", type=int, default=0)
    parser.add_argument("--batch", type=int, default=256)
    parser.add_argument("--precision", choices=["bf16", "fp32"], default="bf16")
    parser.add_argument("--weights", choices=["pretrained", "random"], default="pretrained")
    parser.add_argument("--seed", type=int, default=0)
    return parser


def load_images(paths, transform, device):
    from PIL import Image
    return torch.stack([transform(Image.open(p).convert("RGB")) for p in paths]).to(device)   # transform runs on the device


def main(argv=None) -> None:
    args = build_parser().parse_args(argv)
    rank, world, local = adist.init()
    if not torch.cuda.is_available():
        raise SystemExit("this CLI runs the attack path on CUDA devices (B200); no CPU fallback")
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    say = print if rank == 0 else (lambda *a, **k: None)
    if args.model_type == "robust" and args.weights != "random":
        raise SystemExit("the RobustBench checkpoint (defense_experiments.py:676) is not available offline; "
                         "use --model_type robust --weights random for the same code path on seeded weights")
    # defense_experiments.py:770-772: the JPEG stage of defend_input (run on the device here, bit-exact with PIL's)
    defenses._DEFENSE_USE_JPEG = bool(args.use_jpeg)
    defenses._DEFENSE_JPEG_QUALITY = int(args.jpeg_quality)
    advb200.set_default_precision(args.precision)
    if args.weights == "random":
        torch.manual_seed(args.seed)
        model = models.resnet50(weights=None).eval()
    else:
        try:
            model = models.resnet50(weights=models.ResNet50_Weights.DEFAULT).eval()
        except Exception as e:
            raise SystemExit(f"could not load the pretrained torchvision checkpoint ({e}); use --weights random")
    nm = advb200.native_model(model.to(device))
    if args.model_type == "robust":
        # defense_experiments.py:764-767: the RobustBench model normalises internally, so the attack functions get
        # mean = 0 / std = 1 (same ResNet-50 architecture, the normalisation fused into conv1 becomes the identity)
        mean, std = torch.zeros(3), torch.ones(3)
    else:
        mean = torch.tensor([0.485, 0.456, 0.406])
        std = torch.tensor([0.229, 0.224, 0.225])
    transform = preprocess.get_transform(device)      # Resize(256) -> CenterCrop(224) -> ToTensor, on the device

    # ---- images: the whole list is known to every rank; each rank materialises only its shard
    if args.synthetic > 0:
        n_total = args.synthetic
        gen = torch.Generator().manual_seed(1234)
        all_x = None
        say(f"synthetic images: {n_total}")
    else:
        if args.image_dir is not None:
            image_dir = Path(args.image_dir)
            if not image_dir.is_dir():
                raise SystemExit(f"image_dir not found: {image_dir}")
            image_paths = sorted([p for p in image_dir.iterdir() if p.suffix.lower() in {".jpg", ".jpeg", ".png"}])
            if not image_paths:
                raise SystemExit(f"no images found in {image_dir}")
        else:
            image_path = Path(args.image)
            if not image_path.is_file():
                raise SystemExit(f"image not found: {image_path}")
            image_paths = [image_path]
        n_total = len(image_paths)
    lo, hi = adist.shard_range(n_total, rank, world)
    if args.synthetic > 0:
        full = torch.rand(n_total, 3, 224, 224, generator=gen)      # same stream on every rank
        x_shard = full[lo:hi].to(device)
        calib = full[:min(100, n_total)]
        del full
    else:
        x_shard = load_images(image_paths[lo:hi], transform, device) if hi > lo else \
            torch.empty(0, 3, 224, 224, device=device)
        calib = None

    # ---- detector threshold: calibrated on rank 0 (the first <= 100 images live there), then broadcast
    if args.calibrate_dir is not None:
        calib_dir = Path(args.calibrate_dir)
        if not calib_dir.is_dir():
            raise SystemExit(f"calibrate_dir not found: {calib_dir}")
        calib_paths = sorted([p for p in calib_dir.iterdir() if p.suffix.lower() in {".jpg", ".jpeg", ".png"}])
        if not calib_paths:
            raise SystemExit(f"no images found in calibrate_dir: {calib_dir}")
        thr = defenses.calibrate_feature_threshold(nm, calib_paths, mean=mean, std=std, n=int(args.calibrate_n),
                                                   quantile=float(args.calibrate_quantile), device=device,
                                                   transform=transform) if rank == 0 else 0.0
    elif args.detector_threshold is not None:
        thr = float(args.detector_threshold)
    else:
        if rank == 0:
            src = calib.to(device) if calib is not None else load_images(image_paths[:min(100, n_total)], transform, device)
            thr = defenses.calibrate_feature_threshold(nm, src, mean=mean, std=std, n=min(100, n_total),
                                                       quantile=float(args.calibrate_quantile))
        else:
            thr = 0.0
    thr = adist.broadcast_scalar(thr, src=0, device=device)
    say(f"detector threshold: {thr:.4f}")

    results = {}
    for attack_name in args.attacks:
        for eps in args.eps_list:
            counters = torch.zeros(7, dtype=torch.int64, device=device)
            for b0 in range(0, x_shard.shape[0], args.batch):
                xb = x_shard[b0:b0 + args.batch]
                yb = nm.forward(xb, mean, std, want_logits=False, want_preds=True)["preds"]
                noise = None
                if attack_name == "pgd":
                    # PGD's random start (attack.py:51), drawn per batch from a generator seeded by the batch's GLOBAL
                    # image offset: the noise of an image does not depend on how the list is sharded, so the counters of
                    # a 1-rank and an 8-rank run are bit-equal whenever the shards are whole batches
                    gen = torch.Generator(device=device)
                    gen.manual_seed(int(args.seed) * 1000003 + (lo + b0) * 7 + 1)
                    noise = torch.empty_like(xb).uniform_(-float(eps), float(eps), generator=gen)
                counters += defenses.evaluate_defenses_batched(
                    nm, xb, yb, mean, std, attack_name, float(eps), float(args.alpha), int(args.steps),
                    float(args.cw_c), float(args.cw_kappa), int(args.cw_steps), float(args.cw_lr), thr, pgd_noise=noise)
            adist.allreduce_counters(counters)
            results[(attack_name, float(eps))] = counters.cpu().tolist()

    say("\n" + "=" * 60)
    for (attack_name, eps), c in sorted(results.items()):
        say(f"counters attack={attack_name} eps={eps:.5f}: {c}")
        r = adist.rates_from_counters(c)
        say(f"attack={attack_name}, eps={eps:.5f}, "
            f"attack_success={r['attack_success']:.3f}, "
            f"preproc_defense_acc={r['preproc_defense_acc']:.3f}, "
            f"detector_clean_pass_rate={r['detector_clean_pass_rate']:.3f}, "
            f"detector_adv_flag_rate={r['detector_adv_flag_rate']:.3f}, "
            f"detector_attack_success={r['detector_attack_success']:.3f}")


if __name__ == "__main__":
    main()
