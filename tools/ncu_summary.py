"""Summarise an .ncu-rep (read on the CPU box with `ncu -i … --page raw --csv`) into the few metrics the roofline
discussion needs.  Usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/<name>.txt"""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active", "sm__inst_executed_pipe_tensor", "sm__pipe_tensor_op", "sm__throughput.avg.pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__occupancy_limit",
        "smsp__inst_executed.sum", "sm__cycles_elapsed.max", "sm__cycles_active.avg", "lts__t_sector_hit_rate.pct",
        "lts__t_bytes.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__pcsamp_warps_issue_stalled",
        "sm__inst_executed_pipe_xu", "smsp__issue_active.avg.pct", "smsp__inst_executed_pipe_uniform",
        "launch__shared_mem_per_block_dynamic", "smsp__average_warp"]


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for row in rows[2:]:
        print(f"== launch {row[0]}: {row[4][:100]}  grid {row[8]}  block {row[7]}")
        for i, h in enumerate(hdr):
            if any(w in h for w in WANT) and row[i] not in ("", "n/a"):
                print(f"   {h} = {row[i]} {units[i]}")


if __name__ == "__main__":
    main(sys.argv[1])
