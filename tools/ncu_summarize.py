"""Summaries of ncu CSV exports for profiles/ (run here, on files brought back in gpurun_out/):
    python tools/ncu_summarize.py launches <launch-list.csv> <out.csv>      per-kernel-family time shares of a launch list
    python tools/ncu_summarize.py conv <raw.csv> <out.csv> <traffic.json>   per-launch table + DRAM traffic of the conv family
    python tools/ncu_summarize.py kernels <raw.csv> <out.csv> <out.json>    one row per launch of the HBM-bound update / CW /
                                                                            defense kernels (tools/defense_pass.py capture)"""
import csv
import json
import re
import sys


def family(name):
    for key in ("conv_gemm_kernel", "fused_block_kernel", "prep_hcol", "maxpool_fwd", "maxpool_bwd", "avgpool_fwd", "avgpool_bwd",
                "fc_nt_splitk", "fc_reduce", "loss_grad", "linf_step", "random_start", "argmax"):
        if key in name:
            return key
    return re.sub(r"\(.*", "", name)[:60]


def launches(src, dst):
    rows = [r for r in csv.reader(open(src)) if len(r) > 5]
    hdr = rows[0]
    ki, mi, vi = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
    ui = hdr.index("Metric Unit")
    tot, fam = 0.0, {}
    n = 0
    for r in rows[1:]:
        if r[mi] != "gpu__time_duration.sum":
            continue
        v = float(r[vi].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}.get(r[ui], 1.0)
        f = family(r[ki])
        fam.setdefault(f, [0, 0.0])
        fam[f][0] += 1; fam[f][1] += v
        tot += v; n += 1
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel_family", "launches", "total_us", "share_of_device_time"])
        for k, (c, t) in sorted(fam.items(), key=lambda kv: -kv[1][1]):
            w.writerow([k, c, f"{t:.1f}", f"{t / tot:.4f}"])
        w.writerow(["TOTAL", n, f"{tot:.1f}", "1.0000"])
    conv = sum(t for k, (c, t) in fam.items() if k in ("conv_gemm_kernel", "fused_block_kernel"))
    print(f"{n} launches, {tot / 1e3:.2f} ms of kernel time (cold-cache, serialised by ncu); conv family share {conv / tot:.4f}")


def conv(src, dst, traffic_json):
    rows = list(csv.reader(open(src)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    keep = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32.sum.pct_of_peak_sustained_elapsed" if False else None,
            "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
            "lts__t_sectors_srcunit_tex_op_read.sum", "l1tex__m_xbar2l1tex_read_bytes.sum", "smsp__inst_executed.sum",
            "sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed"]
    keep = [k for k in keep if k and k in idx]
    tensor_cols = [h for h in hdr if "utchmma" in h or "tensor" in h and "pct_of_peak" in h][:4]
    cols = keep + [c for c in tensor_cols if c not in keep]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    tot = 0.0
    n = 0
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(cols); w.writerow([units[idx[c]] for c in cols])
        for r in rows[2:]:
            if "conv_gemm" not in r[idx["Kernel Name"]] and "fused_block" not in r[idx["Kernel Name"]]:
                continue
            w.writerow([r[idx[c]][:80] for c in cols])
            for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                tot += float(r[idx[m]].replace(",", "")) * scale.get(units[idx[m]], 1.0)
            n += 1
    json.dump({"source": f"{dst} (ncu --set full --clock-control none, one forward + input-gradient pass at batch 256; caches flushed per launch by ncu)",
               "launches": n, "dram_bytes_read_plus_write_total": tot, "dram_bytes_per_launch": tot / max(1, n)}, open(traffic_json, "w"), indent=1)
    print(f"{n} conv launches, DRAM read+write {tot / 1e9:.2f} GB per pass, {tot / max(1, n) / 1e6:.1f} MB per launch")
    tcol = "sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed"
    if tcol in idx:
        sel = [r for r in rows[2:] if "conv_gemm" in r[idx["Kernel Name"]] or "fused_block" in r[idx["Kernel Name"]]]
        tt = sum(float(r[idx["gpu__time_duration.sum"]]) for r in sel)
        print("time-weighted bf16 tensor-op (UTCHMMA) rate: %.1f %% of peak over %.0f us" % (
            sum(float(r[idx["gpu__time_duration.sum"]]) * float(r[idx[tcol]]) for r in sel) / tt, tt))


KERNEL_COLS = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
               "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
               "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "smsp__inst_executed.sum",
               "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
               "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
               "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
               "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]


def kernels(src, dst_csv, dst_json):
    rows = list(csv.reader(open(src)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    cols = [c for c in KERNEL_COLS if c in idx]
    out = []
    with open(dst_csv, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(cols); w.writerow([units[idx[c]] for c in cols])
        for r in rows[2:]:
            name = re.sub(r"^void ", "", r[idx["Kernel Name"]]).replace("advb200::<unnamed>::", "").replace("<unnamed>::", "unnamed>::")
            w.writerow([r[idx[c]] for c in cols])
            out.append({"name": re.sub(r"\(.*", "", name).replace("unnamed>::", ""),
                        **{f"{c} [{units[idx[c]]}]": r[idx[c]] for c in cols[1:]}})
    json.dump({"how": "ncu --set full --clock-control none, one warm launch per kernel at N=256 (tools/defense_pass.py, "
                      "tools/gpu_defenses.sh); per-launch values", "kernels": out}, open(dst_json, "w"), indent=1)
    print(f"{len(out)} launches")


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3])
    elif sys.argv[1] == "kernels":
        kernels(sys.argv[2], sys.argv[3], sys.argv[4])
    else:
        conv(sys.argv[2], sys.argv[3], sys.argv[4])
