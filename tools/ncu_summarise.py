#!/usr/bin/env python
"""Distil `ncu -i X.ncu-rep --page raw --csv` into the two small files kept under profiles/:
   <tag>_kernels_full_set.csv  one row per captured kernel, the columns the DESIGN/README readings quote
   <tag>_traffic.json          dram__bytes_read.sum + dram__bytes_write.sum per launch, per kernel (bench.py copies
                               the dominant kernel's entry into roofline.traffic)
   ncu_counts.json             (profiles/) per-launch instruction / DRAM / pipe counters keyed by bench stage name
usage: tools/ncu_summarise.py gpurun_out/<tag>_full_raw.csv profiles/<tag> "<what was captured>" [bench line of the plain run]
"""
import csv
import json
import re
import sys

COLS = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "lts__t_sectors.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio"]
TO_BYTES = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main(raw, out_prefix, what, bench_line=None, counts_out="profiles/ncu_counts.json"):
    rows = list(csv.reader(open(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    cols = [c for c in COLS if c in hdr]
    idx = [hdr.index(c) for c in cols]
    with open(out_prefix + "_kernels_full_set.csv", "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(cols)
        w.writerow([units[i] for i in idx])
        for r in data:
            w.writerow([r[i] for i in idx])
    traffic = {}
    ir, iw, iname = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("Kernel Name")
    for r in data:
        name = re.sub(r"^(void )?(gsr::)?", "", r[iname])
        name = re.split(r"[<(]", name)[0]
        b = float(r[ir]) * TO_BYTES[units[ir]] + float(r[iw]) * TO_BYTES[units[iw]]
        e = traffic.setdefault(name, {"dram_bytes_per_launch": 0.0, "launches_captured": 0})
        e["dram_bytes_per_launch"] += b
        e["launches_captured"] += 1
    for e in traffic.values():
        e["dram_bytes_per_launch"] /= e["launches_captured"]
    traffic["_source"] = what
    json.dump(traffic, open(out_prefix + "_traffic.json", "w"), indent=1)
    # profiles/ncu_counts.json: the per-launch counters bench.py reads (stage name -> instructions, DRAM bytes, pipe
    # utilisation of ONE launch on view 0 of C3; pass the bench line of the plain run as a 4th argument to record the
    # algorithmic bytes of that view beside the measured traffic)
    stage_of = {"preprocess_kernel": "preprocess", "tile_scan_kernel": "scan", "tile_scatter_kernel": "duplicate",
                "tile_sort_kernel": "sort", "render_forward_kernel": "render_forward",
                "render_backward_kernel": "render_backward", "preprocess_backward_kernel": "preprocess_backward"}
    col = lambda r, c: float(r[hdr.index(c)]) if c in hdr and r[hdr.index(c)] not in ("", "n/a") else None
    counts = {}
    for r in data:
        name = re.split(r"[<(]", re.sub(r"^(void )?(gsr::)?", "", r[iname]))[0]
        if name not in stage_of or stage_of[name] in counts:
            continue
        counts[stage_of[name]] = {
            "kernel": name, "inst": col(r, "smsp__inst_executed.sum"),
            "dram_bytes": float(r[ir]) * TO_BYTES[units[ir]] + float(r[iw]) * TO_BYTES[units[iw]],
            "duration_us_under_ncu": col(r, "gpu__time_duration.sum"),
            "issue_active_pct": col(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
            "warps_active_pct": col(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
            "fma_pipe_cycles_pct": col(r, "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
            "alu_pipe_pct": col(r, "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
            "lsu_pipe_pct": col(r, "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
            "shared_wavefronts": col(r, "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"),
            "registers": col(r, "launch__registers_per_thread")}
    if bench_line:
        try:
            line = json.load(open(bench_line))
            for st in line.get("stages", []):
                if st["stage"] in counts:
                    counts[st["stage"]]["algorithmic_bytes_view0"] = st["algorithmic_bytes"]
        except Exception as e:
            print("bench line not used:", e)
    json.dump({"_source": what, "kernels": counts}, open(counts_out, "w"), indent=1)
    for k, e in traffic.items():
        if k != "_source":
            print(f"{k:32s} {e['dram_bytes_per_launch'] / 1e6:9.1f} MB/launch x{e['launches_captured']}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "", sys.argv[4] if len(sys.argv) > 4 else None)
