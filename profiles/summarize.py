"""Turn gpurun_out/*.ncu-rep (ncu --set full captures) — or the raw metric page already exported on the GPU box as
<name>.raw.csv (`ncu -i x.ncu-rep --page raw --csv`) — into the text summaries committed here.
usage: python profiles/summarize.py gpurun_out/r1_add_bf16.ncu-rep gpurun_out/r2_x.raw.csv [...]   -> profiles/<name>.txt"""
import csv
import io
import os
import subprocess
import sys

WANT = [
    "Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes.sum.per_second",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_drain_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_selected_per_issue_active.ratio",
    "launch__cluster_size", "launch__occupancy_limit_warps", "launch__occupancy_limit_blocks", "sm__maximum_warps_per_active_cycle_pct",
    "launch__waves_per_multiprocessor", "sm__ctas_launched.sum",
]

here = os.path.dirname(os.path.abspath(__file__))
for rep in sys.argv[1:]:
    if rep.endswith(".csv"):
        raw = open(rep).read()
        name = os.path.basename(rep)[:-len(".raw.csv")] if rep.endswith(".raw.csv") else os.path.splitext(os.path.basename(rep))[0]
    else:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        name = os.path.splitext(os.path.basename(rep))[0]
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    with open(os.path.join(here, name + ".txt"), "w") as f:
        f.write(f"# ncu --set full --clock-control none, from {os.path.basename(rep)} (one launch per block below)\n")
        for vals in rows[2:]:
            for w in WANT:
                if w in hdr:
                    i = hdr.index(w)
                    f.write(f"{w:84s} {vals[i]:>22s} {units[i]}\n")
            f.write("\n")
    print("wrote", name + ".txt")
