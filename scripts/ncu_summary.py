"""Print the metrics we care about from an .ncu-rep (run here, no GPU needed)."""
import csv, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
keys = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__t_sectors_srcunit_tex_op_read.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_tensor.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_op_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__cycles_elapsed.avg.per_second", "sm__cycles_active.avg", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed_op_shared_atom.sum", "lts__t_sector_hit_rate.pct", "sm__inst_executed_pipe_lsu.sum",
        "smsp__warps_issue_stalled_lg_throttle_per_warp_active.pct", "smsp__warps_issue_stalled_mio_throttle_per_warp_active.pct",
        "smsp__warps_issue_stalled_short_scoreboard_per_warp_active.pct", "smsp__warps_issue_stalled_long_scoreboard_per_warp_active.pct",
        "smsp__warps_issue_stalled_barrier_per_warp_active.pct", "smsp__warps_issue_stalled_membar_per_warp_active.pct"]
pat = sys.argv[2:] 
for r in rows[2:]:
    print("-----")
    for i, h in enumerate(hdr):
        if h in keys or any(p in h for p in pat):
            print(f"{h:90s} {r[i]:>20s} {units[i]}")
