#!/usr/bin/env python
"""Text summary of one .ncu-rep (read in the build container): the headline metrics of B200_PROFILING.md plus the
source lines that execute the most instructions.  python tools/ncu_summary.py gpurun_out/x.ncu-rep > profiles/ncu_x_r02.txt"""
import csv
import io
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_tc_wavefronts_mem_shared.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__inst_executed_pipe_tensor_op_hmma.sum",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tc.sum", "smsp__inst_executed_pipe_fma.sum", "smsp__inst_executed_pipe_alu.sum",
        "smsp__inst_executed_pipe_lsu.sum"]


def run(args):
    return subprocess.run(["ncu", "-i", *args], capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    rows = list(csv.reader(io.StringIO(run([rep, "--page", "raw", "--csv"]))))
    hdr, units, vals = rows[0], rows[1], rows[2]
    name = vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
    print(f"# {rep}\n# kernel: {name}\n# ncu --set full --clock-control none (per-launch, cold cache, serialised)\n")
    for i, h in enumerate(hdr):
        if h in WANT or ("issue_stalled" in h and "per_issue_active" in h and vals[i] and float(vals[i]) >= 0.3):
            print(f"{h:90s} {vals[i]:>16s} {units[i]}")
    src = list(csv.reader(io.StringIO(run([rep, "--page", "source", "--csv", "--print-source", "cuda,sass"]))))
    cur, h2, agg = None, None, []
    for r in src:
        if not r:
            continue
        if r[0] == "File Path":
            cur, h2 = r[1].split("/")[-1], None
            continue
        if r[0] == "Line No":
            h2 = r
            ci, si = h2.index("Instructions Executed"), h2.index("# Samples")
            continue
        if h2 is None or r[0] == "Function Name" or r[2] != "-":
            continue
        try:
            agg.append((float(r[ci]), float(r[si]), cur, r[0], r[1].strip()[:120]))
        except ValueError:
            pass
    tot, ts = sum(a[0] for a in agg) or 1.0, sum(a[1] for a in agg) or 1.0
    print("\n# source lines by executed instructions (share of instructions, share of stall samples)")
    for v, s, f, ln, text in sorted(agg, reverse=True)[:25]:
        print(f"{v / tot * 100:5.1f}% {s / ts * 100:5.1f}%  {f}:{ln:>4}  {text}")


if __name__ == "__main__":
    main()
