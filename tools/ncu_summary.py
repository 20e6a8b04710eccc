#!/usr/bin/env python
"""Condense an `ncu --set full` report into the per-launch table kept under profiles/.

    python tools/ncu_summary.py gpurun_out/X.ncu-rep "header comment" > profiles/X_summary.csv

One column per captured launch, one row per metric (the ones DESIGN.md and bench.py's `roofline`
quote).  Reads the report with `ncu -i ... --page raw --csv`; runs where ncu is installed, no GPU needed.
"""
import csv
import io
import subprocess
import sys

METRICS = [
    "Kernel Name", "Block Size", "Grid Size",
    "gpu__time_duration.sum",
    "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64_op_dmma.sum", "sm__inst_executed_pipe_fp64_op_dmma.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed_op_dmma.sum",
    # fp64 tensor sub-pipe (DMMA) busy cycles per SM sub-partition: divide by smsp__cycles_active.avg for the
    # "DMMA pipe x % of its cycles active" figure DESIGN.md quotes (listed as SM_C.TriageCompute.<name>)
    "smsp__pipe_tensor_subpipe_dmma_cycles_active.avg",
    "sm__inst_executed_pipe_fp64.sum.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "derived__memory_l1_wavefronts_shared_excessive",
    "smsp__thread_inst_executed_per_inst_executed.ratio",
    "sm__sass_thread_inst_executed_op_dfma_pred_on.sum", "sm__sass_thread_inst_executed_op_dmul_pred_on.sum",
    "sm__sass_thread_inst_executed_op_dadd_pred_on.sum",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__cycles_active.avg", "sm__cycles_elapsed.max",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
]


def main():
    rep = sys.argv[1]
    note = sys.argv[2] if len(sys.argv) > 2 else ""
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, launches = rows[0], rows[1], rows[2:]
    idx = {n: i for i, n in enumerate(hdr)}
    for n, i in list(idx.items()):           # "SM_C.TriageCompute.<metric>" -> also reachable as "<metric>"
        if ".Triage" in n:
            idx.setdefault(n.split(".", 2)[2], i)
    out = csv.writer(sys.stdout, lineterminator="\n")
    if note:
        print("# " + note)
    out.writerow(["metric", "unit"] + [f"launch{i}" for i in range(len(launches))])
    for m in METRICS:
        if m not in idx:
            continue
        out.writerow([m, units[idx[m]]] + [r[idx[m]] for r in launches])


if __name__ == "__main__":
    main()
