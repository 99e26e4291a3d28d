#!/usr/bin/env python
"""Condense ncu outputs from gpurun_out/ into the tracked summaries under profiles/.

    python profiles/summarize.py launches gpurun_out/launches_r1.csv profiles/r1_launches_summary.txt
    python profiles/summarize.py full gpurun_out/prof_edge_r1.ncu-rep profiles/r1_edge_pass_full.txt [traffic.json]
"""
import collections
import csv
import io
import json
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__occupancy_limit_registers",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "sm__cycles_elapsed.max",
    "sm__inst_executed_pipe_tensor.sum", "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
    "lts__t_bytes.sum", "lts__t_sectors_srcunit_tex_op_read.sum", "l1tex__m_xbar2l1tex_read_bytes.sum",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "launch__shared_mem_per_block_dynamic",
    "smsp__cycles_active.avg", "sm__cycles_active.avg",
]
OURS = ("persistent_kernel", "persistent_batch_kernel", "persistent_small_kernel", "gram_kernel", "gram_split_kernel", "gram_colmax_kernel", "gram_norm_reduce_kernel",
        "edge_pass_kernel", "node_tail_kernel", "edge_init_kernel", "theta_rows_kernel", "theta_mean_kernel",
        "pdist_kernel", "knn_", "pack_to_tiles", "tiles_to_", "store_packed", "gram_edges_kernel")
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}


def launches(src, dst):
    rows = [l for l in open(src) if l.startswith('"')]
    recs = list(csv.DictReader(io.StringIO("".join(rows))))
    agg = collections.defaultdict(lambda: [0, 0.0])
    for x in recs:
        v = float(x["Metric Value"].replace(",", "")) * UNIT.get(x["Metric Unit"], 1.0)  # -> us
        agg[x["Kernel Name"][:90]][0] += 1
        agg[x["Kernel Name"][:90]][1] += v
    tot = sum(v[1] for v in agg.values())
    is_ours = lambda k: "gl::" in k or any(t in k for t in OURS)  # noqa: E731  (newer ncu drops the namespace)
    ours = sum(v[1] for k, v in agg.items() if is_ours(k))
    with open(dst, "w") as f:
        f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none: {len(recs)} launches, "
                f"{tot / 1e3:.3f} ms total, {ours / 1e3:.3f} ms in libgl_b200 kernels (gl::)\n")
        f.write("# per-launch times are cold-cache and serialised: compare SHARES\n")
        f.write(f"{'kernel':92s} {'count':>6s} {'total_ms':>10s} {'avg_us':>10s} {'share':>7s} {'share_of_gl':>11s}\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            sg = f"{v[1] / ours:11.3f}" if is_ours(k) else " " * 11
            f.write(f"{k:92s} {v[0]:6d} {v[1] / 1e3:10.3f} {v[1] / v[0]:10.1f} {v[1] / tot:7.3f} {sg}\n")


def full(src, dst, traffic=None):
    out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(out)))
    h, units, data = r[0], r[1], r[2:]
    with open(dst, "w") as f:
        f.write(f"# ncu --set full --clock-control none --import-source on; source: {src}; {len(data)} launches\n")
        kn = h.index("Kernel Name")
        f.write("kernel: " + data[0][kn] + "\n")
        for m in METRICS:
            if m in h:
                i = h.index(m)
                f.write(f"{m:86s} [{units[i]}] " + "  ".join(x[i] for x in data) + "\n")
    if traffic:
        ir, iw = h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum")
        vals = [float(x[ir]) * UNIT[units[ir]] + float(x[iw]) * UNIT[units[iw]] for x in data]
        with open(traffic, "w") as f:
            json.dump({"dram_bytes_per_launch": sum(vals) / len(vals), "launches": len(vals), "source": src}, f)


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3])
    else:
        full(sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else None)
