#!/usr/bin/env python
"""Text summary of one kernel of an .ncu-rep (the files under profiles/ are made with this).

    python scripts/ncu_summary.py gpurun_out/prof_r02.ncu-rep tc2_res_kernel > profiles/r02_ncu_full_tc2_res_kernel.txt

Prints the headline metrics of the `--set full` capture (duration, pipe utilisation, DRAM/L2 traffic,
occupancy), the warp-stall totals from the source page, and the hottest SASS instructions with their
dominant stall reason.  Needs `ncu` on PATH (reads the report; no GPU).
"""
import csv
import io
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.max", "launch__grid_size", "launch__block_size",
    "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
]


def ncu_csv(rep, page, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv", *extra], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep, kernel = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    raw = ncu_csv(rep, "raw")
    hdr, units = raw[0], raw[1]
    col = {h: i for i, h in enumerate(hdr)}
    row = next(r for r in raw[2:] if kernel in r[col["Kernel Name"]])
    print("# %s  (%s)" % (row[col["Kernel Name"]], rep))
    print("# ncu --set full --import-source on --clock-control none; one launch; times under ncu are cold-cache/serialised")
    for m in METRICS:
        if m in col:
            print("%-70s %14s %s" % (m, row[col[m]], units[col[m]]))
    src = ncu_csv(rep, "source", ["--print-source", "sass", "--kernel-name", "regex:" + kernel])
    shdr = src[1]
    sc = {h: i for i, h in enumerate(shdr)}
    data = [r for r in src[2:] if len(r) == len(shdr)]
    S, E = sc["Warp Stall Sampling (All Samples)"], sc["Instructions Executed"]
    stalls = [h for h in shdr if h.startswith("stall_") and "Not Issued" not in h]
    total = sum(int(r[S]) for r in data)
    print("\n# warp stall samples (all warps): %d" % total)
    agg = sorted(((sum(int(r[sc[h]]) for r in data), h) for h in stalls), reverse=True)
    for v, h in agg:
        if v:
            print("%-28s %9d  %5.1f %%" % (h, v, 100.0 * v / total))
    print("\n# hottest SASS instructions (index, samples, share, executed, top stall reasons)")
    hot = sorted(range(len(data)), key=lambda i: -int(data[i][S]))[:top]
    for i in sorted(hot):
        r = data[i]
        s = int(r[S])
        why = sorted(((int(r[sc[h]]), h[6:]) for h in stalls), reverse=True)[:2]
        print("%5d %-72s %8d %5.1f %% %12s  %s" % (i, r[1].strip()[:72], s, 100.0 * s / total, r[E],
                                                   ", ".join("%s %d" % (n, v) for v, n in why if v)))


if __name__ == "__main__":
    main()
