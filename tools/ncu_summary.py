"""Summarise ncu outputs under gpurun_out/ into small text files under profiles/ (tracked).

  python tools/ncu_summary.py launches gpurun_out/launches.csv profiles/r01_launches.md "title"
  python tools/ncu_summary.py raw gpurun_out/prof_x.ncu-rep|prof_x_raw.csv profiles/r01_x_raw.md "title"
"""
import collections
import csv
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__waves_per_multiprocessor", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__cycles_elapsed.max",
    "l1tex__t_sector_pipe_lsu_mem_local_op_ld_hit_rate.pct", "l1tex__t_sector_pipe_lsu_mem_local_op_st_hit_rate.pct",
    "sass__inst_executed_local_loads", "sass__inst_executed_local_stores",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "sm__icc_request_hit_rate.pct",
]


def launches(src, dst, title):
    lines = [l for l in open(src) if l.startswith('"')]
    r = csv.reader(lines)
    hdr = next(r)
    ki, vi, ui, gi, bi = (hdr.index(k) for k in ("Kernel Name", "Metric Value", "Metric Unit", "Grid Size", "Block Size"))
    agg = collections.OrderedDict()
    tot, n = 0.0, 0
    for row in r:
        v = float(row[vi].replace(",", ""))
        v = {"ns": v / 1e3, "us": v, "usecond": v, "ms": v * 1e3, "msecond": v * 1e3}.get(row[ui], v)
        k = row[ki].split("(")[0]
        a = agg.setdefault(k, [0, 0.0, 0.0])
        a[0] += 1
        a[1] += v
        a[2] = max(a[2], v)
        tot += v
        n += 1
    with open(dst, "w") as f:
        f.write("# %s\n\n" % title)
        f.write("Source: `ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off` "
                "(cold-cache, serialised launches: compare SHARES, not absolutes).\n\n")
        f.write("%d launches, %.1f us total kernel time\n\n" % (n, tot))
        f.write("| kernel | launches | total us | share | max us |\n|---|---|---|---|---|\n")
        for k, (c, t, m) in sorted(agg.items(), key=lambda x: -x[1][1]):
            f.write("| %s | %d | %.1f | %.1f%% | %.1f |\n" % (k, c, t, 100 * t / tot, m))


def raw(src, dst, title):
    if src.endswith(".csv"):   # exported on the GPU box: ncu -i x.ncu-rep --page raw --csv > x_raw.csv
        out = open(src).read()
    else:
        out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ki = hdr.index("Kernel Name")
    with open(dst, "w") as f:
        f.write("# %s\n\nSource: `ncu --set full --clock-control none --import-source on` (%s), %d launches captured.\n\n"
                % (title, src, len(data)))
        f.write("| metric | unit | " + " | ".join("launch %d" % i for i in range(len(data))) + " |\n")
        f.write("|---|---|" + "---|" * len(data) + "\n")
        f.write("| kernel | | " + " | ".join(d[ki].split("(")[0] for d in data) + " |\n")
        for k in KEYS:
            for i, h in enumerate(hdr):
                if h == k:
                    f.write("| %s | %s | %s |\n" % (h, units[i], " | ".join(d[i] for d in data)))


if __name__ == "__main__":
    {"launches": launches, "raw": raw}[sys.argv[1]](sys.argv[2], sys.argv[3], sys.argv[4])
