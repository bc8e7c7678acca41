"""Print selected metrics from `ncu -i X.ncu-rep --page raw --csv` (one column per captured launch)."""
import csv
import subprocess
import sys

WANT = [
    "Kernel Name", "gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "lts__t_bytes.sum", "sm__pipe_tensor_cycles_active", "sm__inst_executed_pipe_tensor",
    "smsp__issue_active.avg.pct", "smsp__inst_executed.sum",
    "issue_stalled_barrier_per_warp_active", "issue_stalled_long_scoreboard_per_warp_active",
    "issue_stalled_short_scoreboard_per_warp_active", "issue_stalled_math_pipe_throttle_per_warp_active",
    "issue_stalled_mio_throttle_per_warp_active", "issue_stalled_lg_throttle_per_warp_active",
    "issue_stalled_wait_per_warp_active", "issue_stalled_membar_per_warp_active",
    "issue_stalled_sleeping_per_warp_active", "issue_stalled_not_selected_per_warp_active",
    "issue_stalled_dispatch_stall", "bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
]


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    for w in WANT:
        for i, h in enumerate(hdr):
            if w in h:
                vals = [r[i][:48] for r in data]
                print(f"{h} [{units[i]}]: {vals}")


if __name__ == "__main__":
    main(sys.argv[1])
