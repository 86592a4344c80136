#!/usr/bin/env python
"""bench.py — PGD-10 adversarial images/sec, ResNet-50 224^2 bf16 (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU path (oracle port)

A step = one PGD-10 attack (eps=8/255, alpha=2/255, random start) over 256 synthetic images per GPU
(config "ResNet-50 PGD-10 ... batch=256 bf16 on 1 B200"); weak scaling: every rank attacks its own
256-image batch.  `value` is timed with the images resident in HBM; `e2e` goes through the public
Python API from pinned host buffers (H2D of the images, D2H of the adversarial images) every step.

    python bench.py --config shard8192 --gpus N ...           # BASELINE configs[3]: 8192 images in total, contiguous
                                                              # shards of 8192/N per GPU (N = 1: chunks of 256), the
                                                              # success counters summed by one all-reduce; "strong"
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
PKG = ROOT / "image-recognition-adversarial-example-attack_b200"
for p in (str(ROOT), str(PKG)):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402

BATCH = 256
PGD_STEPS = 10
EPS, ALPHA = 8 / 255, 2 / 255
GFLOP_PER_IMG_STEP = 16.357           # fwd + dgrad, 2 flop/MAC
GFLOP_PER_IMG_STEP_LITERAL = 8.2      # BASELINE.json's wording (= the MAC count)
METRIC = "PGD-10 adv images/sec, ResNet-50 224^2 bf16"


def measured_peaks():
    f = ROOT / "MEASURED_PEAKS.json"
    if f.is_file():
        d = json.loads(f.read_text())
        return d, "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i] == "Active"})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def _reference_attack_module():
    """The reference's own, unmodified attack.py when build() staged it under oracle/_ref/ (oracle/make_ref.py copies the
    file from /root/reference in the build container; the directory is git-ignored and travels with the snapshot)."""
    ref_file = ROOT / "oracle" / "_ref" / "attack.py"
    if not ref_file.is_file():
        return None
    import importlib.util
    spec = importlib.util.spec_from_file_location("reference_attack", ref_file)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


_REF = _reference_attack_module()
REF_KIND = "reference" if _REF is not None else "port"
REF_WHAT = ("the reference's unmodified attack.pgd_linf_attack on torchvision resnet50 (CPU, fp32)" if _REF is not None
            else "oracle port of attack.pgd_linf_attack")


def cpu_port_pgd(n_img: int, threads: int):
    """The reference's CPU path on the host cores — its own attack.pgd_linf_attack when staged (kind "reference"), else
    the oracle port (kind "port"): returns (img/s, seconds)."""
    from oracle import attacks_ref, resnet50_ref, weights
    torch.set_num_threads(threads)
    model = weights.seeded_resnet50(0)
    mean, std = weights.mean_std()
    x = weights.synthetic_images(n_img)
    if _REF is not None:
        with torch.no_grad():
            y = model(_REF.normalize_batch(x, mean, std)).argmax(1)
        run = lambda xs, ys, steps: _REF.pgd_linf_attack(model, xs, ys, eps=EPS, alpha=ALPHA, steps=steps, mean=mean, std=std)
    else:
        fm = resnet50_ref.FunctionalResNet50(model.state_dict())
        with torch.no_grad():
            y = fm(attacks_ref.normalize(x, mean, std)).argmax(1)
        noise = weights.pgd_noise(x.shape, EPS, 1)
        run = lambda xs, ys, steps: attacks_ref.pgd_linf(fm, xs, ys, eps=EPS, alpha=ALPHA, steps=steps, mean=mean, std=std,
                                                         noise=noise[:xs.shape[0]])
    run(x[:2], y[:2], 1)
    t0 = time.perf_counter()
    for lo in range(0, n_img, 16):     # batches of 16: the size at which the CPU path is most efficient (measured)
        run(x[lo:lo + 16], y[lo:lo + 16], PGD_STEPS)
    dt = time.perf_counter() - t0
    return n_img / dt, dt


def host_threads() -> int:
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = host_threads()
    n_img = 16
    vals = []
    for _ in range(args.warmup_ref):
        cpu_port_pgd(2, threads)
    t_all = time.perf_counter()
    for _ in range(args.steps_ref):
        v, _ = cpu_port_pgd(n_img, threads)
        vals.append(v)
    wall = time.perf_counter() - t_all
    val = sorted(vals)[len(vals) // 2]
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "images/s", "n_gpus": args.gpus,
        "steps": args.steps_ref, "warmup": args.warmup_ref, "ms_per_step": 1e3 * wall / max(1, args.steps_ref),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "ResNet-50 (seed-0 random init) PGD-10 eps=8/255 alpha=2/255, synthetic 224x224",
                   "batch_per_step": n_img, "note": "reference's CPU PyTorch path: " + REF_WHAT},
        "cpu_baseline": {"value": val, "unit": "images/s", "cores": threads, "kind": REF_KIND,
                         "sample": f"PGD-10 on {n_img} images per step (bounded sample of the 256-image batch)"},
        "e2e": {"value": val, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["native", "reference"], default="native")
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--config", choices=["pgd256", "shard8192"], default="pgd256",
                    help="pgd256: the headline (256 images per GPU, weak); shard8192: 8192 images sharded over the GPUs (strong)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.steps_ref = max(1, min(args.steps, 3))
    args.warmup_ref = max(1, min(args.warmup, 1))
    if args.impl == "reference":
        run_reference(args)
        return

    import advb200
    from advb200 import _lib, dist as adist
    from oracle import weights          # seeded weights / images only (shared test inputs)

    rank, world, local = adist.init()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the attack hot path has no CPU fallback")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    warm = max(3, args.warmup)
    n = args.batch
    strong = args.config == "shard8192"
    if strong:
        total = 8192
        lo, hi = adist.shard_range(total, rank, world)      # contiguous shard of this rank
        n = hi - lo

    model = weights.seeded_resnet50(0).to(dev)
    mean, std = weights.mean_std()
    nm = advb200.NativeResNet50(model, precision="bf16")
    if strong:
        # every rank draws its own shard's images (seeded per 256-image chunk, so the global image list is the same for
        # every N); the attack runs chunk by chunk inside pgd_linf_attack (ADVB200_MAX_CHUNK = 256)
        x = torch.cat([weights.synthetic_images(256, seed=1234 + c) for c in range(lo // 256, (hi + 255) // 256)])
        x = x[lo - (lo // 256) * 256:][:n].to(dev)
    else:
        x = weights.synthetic_images(n, seed=1234 + rank).to(dev)
    y = nm.forward(x, mean, std, want_logits=False, want_preds=True)["preds"]

    def attack(xin):
        return advb200.pgd_linf_attack(nm, xin, y, eps=EPS, alpha=ALPHA, steps=PGD_STEPS, mean=mean, std=std)

    for _ in range(warm):
        xa = attack(x)
    torch.cuda.synchronize()

    # ---------------- device-resident timing (value) ----------------
    sampler = ClockSampler(torch.cuda.current_device() if "CUDA_VISIBLE_DEVICES" not in os.environ else local)
    sampler.start()
    l0 = _lib.launch_count()
    adist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    prof_range = bool(os.environ.get("ADVB200_BENCH_PROFILE_RANGE"))    # ncu --profile-from-start off: only the timed steps
    if prof_range:
        torch.cuda.profiler.start()
    e0.record()
    for _ in range(args.steps):
        xa = attack(x)
    e1.record()
    if prof_range:
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    torch.cuda.synchronize(); adist.barrier()
    ms = adist.max_over_ranks(e0.elapsed_time(e1), device=dev)
    launches = _lib.launch_count() - l0
    clocks = sampler.stop()
    n_all = int(adist.allreduce_counters(torch.tensor([n], dtype=torch.int64, device=dev))[0])   # images per step, all ranks
    value = n_all * args.steps / (ms / 1e3)

    # success counters, summed over ranks by one all-reduce (NCCL over NVLink for N > 1)
    pa = nm.forward(xa, mean, std, want_logits=False, want_preds=True)["preds"]
    counters = torch.tensor([int((pa != y).sum()), n], dtype=torch.int64, device=dev)
    adist.allreduce_counters(counters)

    # ---------------- end to end through the public API from pinned host memory ----------------
    # every step: H2D of that step's pinned input batch, the attack, D2H of its result.  The public host-buffer
    # front end (advb200.pipeline.HostBatchPipeline) overlaps the copies of neighbouring steps with the attack.
    import itertools
    from advb200.pipeline import HostBatchPipeline
    chunks_per_step = (n + BATCH - 1) // BATCH if strong else 1
    if strong:        # the shard streams through the pipeline as 256-image pinned chunks, each with its own labels
        xh_list = [x[c * BATCH:(c + 1) * BATCH].cpu().pin_memory() for c in range(chunks_per_step)]
        y_list = [y[c * BATCH:(c + 1) * BATCH] for c in range(chunks_per_step)]
        assert all(t.shape[0] == BATCH for t in xh_list), "shard8192 expects shards that are multiples of 256"
        state = {"i": 0}

        def attack_chunk(xin):
            yy = y_list[state["i"] % chunks_per_step]
            state["i"] += 1
            return advb200.pgd_linf_attack(nm, xin, yy, eps=EPS, alpha=ALPHA, steps=PGD_STEPS, mean=mean, std=std,
                                           inv_n_global=1.0 / 8192)
        pipe = HostBatchPipeline(attack_chunk, dev)
        xh = xh_list[0]
        feed = lambda steps: (xh_list[i % chunks_per_step] for i in range(steps * chunks_per_step))
    else:
        xh = x.cpu().pin_memory()
        pipe = HostBatchPipeline(attack, dev)
        feed = lambda steps: (xh for _ in range(steps))
    outs = [torch.empty_like(xh).pin_memory() for _ in range(2)]
    for _ in pipe.run(feed(2 if not strong else 1), itertools.cycle(outs)):
        pass
    adist.barrier(); torch.cuda.synchronize()
    trace = os.environ.get("ADVB200_E2E_TRACE")
    marks = []
    if trace:                      # debug: device timeline of every step of the end-to-end loop
        inner = pipe.fn

        def traced(xd):
            a = torch.cuda.Event(enable_timing=True); a.record()
            r = inner(xd)
            b = torch.cuda.Event(enable_timing=True); b.record()
            marks.append((a, b, time.perf_counter()))
            return r
        pipe.fn = traced
    t_host0 = time.perf_counter()
    e0.record()
    for _ in pipe.run(feed(args.steps), itertools.cycle(outs)):
        pass
    torch.cuda.current_stream().wait_stream(pipe.d2h)
    e1.record()
    torch.cuda.synchronize(); adist.barrier()
    if trace and rank == 0:
        print("e2e trace: total %.1f ms; attack start %s dur %s host-return %s" % (
            e0.elapsed_time(e1), [round(e0.elapsed_time(m[0]), 1) for m in marks],
            [round(m[0].elapsed_time(m[1]), 1) for m in marks],
            [round((m[2] - t_host0) * 1e3, 1) for m in marks]), file=sys.stderr, flush=True)
    ms_e2e = adist.max_over_ranks(e0.elapsed_time(e1), device=dev)
    e2e = n_all * args.steps / (ms_e2e / 1e3)
    nbytes = xh.numel() * 4 * chunks_per_step

    # ---------------- roofline of the dominant kernel (tcgen05 conv family) ----------------
    lib = _lib.load()
    import ctypes
    lib.advb200_profile_enable(1)
    torch.cuda.synchronize()
    n_prof = min(n, BATCH) if strong else n
    advb200.pgd_linf_attack(nm, x[:n_prof], y[:n_prof], eps=EPS, alpha=ALPHA, steps=PGD_STEPS, mean=mean, std=std)
    torch.cuda.synchronize()
    conv_ms, conv_n = ctypes.c_double(), ctypes.c_int64()
    lib.advb200_profile_read(ctypes.byref(conv_ms), ctypes.byref(conv_n))
    lib.advb200_profile_enable(0)
    peaks, peak_kind = measured_peaks()
    conv_flops = n_prof * PGD_STEPS * GFLOP_PER_IMG_STEP * 1e9     # algorithmic, per attack
    per_launch_flops = conv_flops / max(1, conv_n.value)
    per_launch_ms = conv_ms.value / max(1, conv_n.value)
    achieved = per_launch_flops / (per_launch_ms * 1e-3) / 1e12 if per_launch_ms > 0 else 0.0
    peak = float(peaks.get("bf16_tflops_sustained", peaks["bf16_tflops"]))
    step_tflops = value / world * PGD_STEPS * GFLOP_PER_IMG_STEP / 1e3

    if rank != 0:
        return
    # DRAM traffic of the conv launches from the committed ncu --set full capture (per launch, like `achieved`)
    traffic, traffic_src = None, None
    for tname in ("r02_conv_traffic.json", "r01_conv_traffic.json"):
        tfile = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", tname)
        if os.path.isfile(tfile) and n_prof == BATCH:
            traffic = json.load(open(tfile)).get("dram_bytes_per_launch")
            traffic_src = f"profiles/{tname}: committed ncu --set full capture of the same launches (not measured in this run)"
            break
    cpu = None
    if not args.no_cpu_baseline and world == 1:
        threads = host_threads()
        v, dt = cpu_port_pgd(64, threads)
        cpu = {"value": v, "unit": "images/s", "cores": threads, "kind": REF_KIND,
               "sample": f"{REF_WHAT}, PGD-10 on 64 of the 256 images in batches of 16, {dt:.1f} s"}
    line = {
        "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": warm,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong" if strong else "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": ("ResNet-50 (seed-0 random init) PGD-10 eps=8/255 alpha=2/255, synthetic 224x224, "
                                + (f"batch=8192 sharded over {world} GPU(s) in contiguous shards ({n} per GPU, chunks of 256), "
                                   "success counters by one all-reduce" if strong else f"batch={n} per GPU")),
                   "batch_per_gpu": n, "pgd_steps": PGD_STEPS,
                   "l2": "inputs + activations (GBs per step) exceed the 126 MB L2; no explicit flush",
                   "attack_success": int(counters[0]) / max(1, int(counters[1]))},
        "e2e": {"value": e2e, "unit": "images/s", "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": nbytes},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                     "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                     "kernel": "conv_gemm_kernel<BN> + fused_block_kernel (tcgen05 implicit-GEMM convs, all fwd+dgrad launches)",
                     "peak_source": f"{peak_kind} bf16_tflops_sustained (kernel timed inside a long step)",
                     "launches_per_attack": int(conv_n.value), "conv_ms_per_attack": conv_ms.value,
                     "whole_step_tflops": step_tflops, "whole_step_frac": step_tflops / peak,
                     "achieved_baseline_literal": achieved * GFLOP_PER_IMG_STEP_LITERAL / GFLOP_PER_IMG_STEP},
        "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    try:
        main()
    finally:
        try:
            from advb200 import dist as _adist
            _adist.shutdown()
        except Exception:
            pass
