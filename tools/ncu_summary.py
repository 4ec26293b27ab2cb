#!/usr/bin/env python
"""Summarise ncu outputs brought back in gpurun_out/ into the small text files kept under profiles/.

    python tools/ncu_summary.py launches gpurun_out/launches_r01.csv  > profiles/r01_launches.txt
    python tools/ncu_summary.py report   gpurun_out/prof_x.ncu-rep    > profiles/r01_x_metrics.txt
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__waves_per_multiprocessor", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__cycles_active.avg", "sm__inst_executed_pipe_uniform.sum",
    # shared memory / L2 / tensor pipe: what "smem-bound" and "issue-bound" claims rest on
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "smsp__inst_executed_op_shared_ld.sum", "smsp__inst_executed_op_shared_st.sum",
    "lts__t_sector_hit_rate.pct", "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_write.sum",
    "sm__inst_executed_pipe_tensor.sum", "sm__inst_executed_pipe_tensor_subpipe_hmma.sum",
    "sm__inst_executed_pipe_lsu.sum", "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_fma.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
]
# warp-stall breakdown (WarpStateStats section of --set full): cycles a warp spends stalled for each reason per
# instruction it issues; the summary prints the five largest
STALL_PREFIXES = ("smsp__average_warps_issue_stalled_", "smsp__average_warp_latency_issue_stalled_",
                  "smsp__pcsamp_warps_issue_stalled_")


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    n = 0
    for row in csv.DictReader(io.StringIO("".join(lines))):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        n += 1
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(row["Metric Unit"], 1.0)
        k = row["Kernel Name"].split("(")[0]
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print(f"# {path}: {n} launches, {tot / 1e6:.3f} ms of kernel time (ncu: cold-cache, serialised — compare shares)")
    print(f"{'kernel':60s} {'n':>6s} {'sum_us':>12s} {'avg_us':>10s} {'share':>7s}")
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"{k[:60]:60s} {a[0]:6d} {a[1] / 1e3:12.1f} {a[1] / a[0] / 1e3:10.2f} {a[1] / tot * 100:6.1f}%")


def report(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    print(f"# {path} (ncu --set full --clock-control none)")
    for r in rows[2:]:
        print("kernel:", r[hdr.index("Kernel Name")][:110])
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print(f"    {k:70s} {r[i]:>18s} {units[i]}")
        stalls = []
        for i, h in enumerate(hdr):
            if h.startswith(STALL_PREFIXES) and not h.endswith("_not_issued") and r[i] not in ("", "n/a"):
                try:
                    stalls.append((float(r[i].replace(",", "")), h, units[i]))
                except ValueError:
                    pass
        for fam in STALL_PREFIXES:
            top = sorted((x for x in stalls if x[1].startswith(fam)), reverse=True)[:5]
            for v, h, u in top:
                print(f"    stall  {h:63s} {v:18.3f} {u}")


if __name__ == "__main__":
    {"launches": launches, "report": report}[sys.argv[1]](sys.argv[2])
