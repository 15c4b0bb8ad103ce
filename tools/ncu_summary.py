#!/usr/bin/env python
"""Turn ncu outputs brought back in gpurun_out/ into the small text summaries kept under profiles/.

    python tools/ncu_summary.py launches <launches.csv> <out.md> [--particles N --cells C --pairs P [--peak GBps]]
                                                  # per-kernel totals and shares of a launch list (+ GB/s per kernel)
    python tools/ncu_summary.py full <report.ncu-rep> <out.csv> [--traffic-json profiles/roofline_traffic.json]
"""
import collections
import csv
import io
import json
import subprocess
import sys

SCALE = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3}
KEEP = [
    "Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg",
    "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "smsp__cycles_active.avg",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct",
    "smsp__warp_issue_stalled_lg_throttle_per_warp_active.pct", "smsp__warp_issue_stalled_barrier_per_warp_active.pct",
    "smsp__warp_issue_stalled_membar_per_warp_active.pct",
]
BYTES = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def measured_peak():
    """hbm_gbs of MEASURED_PEAKS.json at the repo root (the one source of the roofline denominator); fallback as B200_PROFILING.md states."""
    import os
    try:
        with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"])
    except Exception:
        return 6650.0


def launches(path, out, algo=None):
    with open(path) as fh:
        lines = [l for l in fh if not l.startswith("==")]
    agg, cnt = collections.OrderedDict(), collections.Counter()
    for r in csv.DictReader(io.StringIO("".join(lines))):
        try:
            v = float(r["Metric Value"].replace(",", "")) * SCALE[r["Metric Unit"]]
        except (KeyError, ValueError):
            continue
        name = r["Kernel Name"].split("(")[0].replace("void ", "")
        agg[name] = agg.get(name, 0.0) + v
        cnt[name] += 1
    tot = sum(agg.values())
    with open(out, "w") as f:
        f.write(f"# ncu launch list summary of `{path}`\n\n")
        f.write("`ncu --metrics gpu__time_duration.sum --clock-control none` - per-launch times are cold-cache and serialised: compare shares, not absolutes.\n\n")
        f.write(f"total device time of the captured launches: {tot:.3f} ms over {sum(cnt.values())} launches\n\n")
        f.write("| kernel | launches | total ms | avg ms | share |\n|---|---:|---:|---:|---:|\n")
        for k, v in sorted(agg.items(), key=lambda x: -x[1]):
            f.write(f"| `{k}` | {cnt[k]} | {v:.3f} | {v / cnt[k]:.4f} | {100 * v / tot:.1f}% |\n")
        if algo:
            # algorithmic bytes of the collision kernels (fp32 records), DESIGN.md section 4: per particle, per cell or per listed pair
            per = {"bin_kernel": ("particle", 24), "place_kernel": ("particle", 16), "move_kernel": ("particle", 48),
                   "scan_reduce_kernel": ("cell", 1), "scan_down_kernel": ("cell", 5), "detect_kernel": ("particle", 20),
                   "resolve_kernel": ("pair", 100), "integrate_bounce_kernel": ("particle", 32)}
            f.write(f"\nPer kernel against the HBM roofline ({algo['particles']} particles, {algo['cells']} cells, {algo['pairs']} listed pairs; "
                    f"peak {algo['peak']} GB/s measured copy bandwidth):\n\n")
            f.write("| kernel | avg ms | particles/s | algorithmic bytes per launch | GB/s | of peak |\n|---|---:|---:|---:|---:|---:|\n")
            for k, v in sorted(agg.items(), key=lambda x: -x[1]):
                base = k.split("<")[0]
                if base not in per:
                    continue
                unit, nbytes = per[base]
                ms = v / cnt[k]
                total = nbytes * algo[{"particle": "particles", "cell": "cells", "pair": "pairs"}[unit]]
                gbs = total / (ms * 1e-3) / 1e9
                f.write(f"| `{k}` | {ms:.4f} | {algo['particles'] / (ms * 1e-3):.3g} | {total / 1e6:.0f} MB ({nbytes} B per {unit}) | {gbs:.0f} | {100 * gbs / algo['peak']:.0f}% |\n")
    print(open(out).read())


def full(rep, out, traffic_json=None):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    cols = [hdr.index(k) for k in KEEP if k in hdr]
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow([hdr[c] for c in cols])
        w.writerow([units[c] for c in cols])
        for r in rows[2:]:
            w.writerow([r[c] for c in cols])
    per = collections.defaultdict(list)
    ir, iw, ik = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("Kernel Name")
    for r in rows[2:]:
        b = float(r[ir]) * BYTES[units[ir]] + float(r[iw]) * BYTES[units[iw]]
        name = r[ik].split("(")[0].replace("void ", "")
        per[name].append(b)
    summary = {k: sum(v) / len(v) for k, v in per.items()}
    print(json.dumps(summary, indent=1))
    if traffic_json:
        try:
            cur = json.load(open(traffic_json))
        except Exception:
            cur = {}
        for k, v in summary.items():
            base = k.split("<")[0]
            targ = k.split("<")[1].split(",")[0].rstrip(">") if "<" in k else ""
            cur[f"{base}<{targ}>" if targ else base] = v
        json.dump(cur, open(traffic_json, "w"), indent=1)


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        algo = None
        if "--particles" in sys.argv:
            def opt(name, default):
                return float(sys.argv[sys.argv.index(name) + 1]) if name in sys.argv else default
            algo = {"particles": int(opt("--particles", 0)), "cells": int(opt("--cells", 0)), "pairs": int(opt("--pairs", 0)),
                    "peak": opt("--peak", measured_peak())}
        launches(sys.argv[2], sys.argv[3], algo)
    else:
        tj = sys.argv[sys.argv.index("--traffic-json") + 1] if "--traffic-json" in sys.argv else None
        full(sys.argv[2], sys.argv[3], tj)
