"""tools/ncu_summarize.py -- read gpurun_out/ncu/*.ncu-rep (here, no GPU needed) and write the metrics the
roofline claims rest on into profiles/<prefix>_<name>_ncu.txt.   usage: python tools/ncu_summarize.py r01"""
import csv
import glob
import io
import os
import subprocess
import sys

KEYS = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__ops_path_tensor_src_fp64.sum.per_second", "sm__ops_path_tensor_src_tf32_dst_fp32.sum.per_second",
        "sm__inst_executed_pipe_uniform.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum", "l1tex__t_sector_hit_rate.pct",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__cycles_elapsed.avg.per_second",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio")


def main():
    prefix = sys.argv[1] if len(sys.argv) > 1 else "r01"
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for rep in sorted(glob.glob(os.path.join(root, "gpurun_out", "ncu", "*.ncu-rep"))):
        name = os.path.basename(rep)[:-8]
        out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(out)))
        if len(rows) < 3:
            print("no data in", rep)
            continue
        hdr, units = rows[0], rows[1]
        lines = [f"# ncu --set full --clock-control none -k <kernel> -c 1 (tools/ncu_capture.sh); report {os.path.basename(rep)}"]
        for r in rows[2:]:
            d = dict(zip(hdr, r))
            lines.append(f"# kernel: {d.get('Kernel Name')}  grid {d.get('Grid Size')} block {d.get('Block Size')}")
            for k in KEYS:
                if k in d and d[k] != "":
                    lines.append(f"{k:90s} {units[hdr.index(k)]:16s} {d[k]}")
        path = os.path.join(root, "profiles", f"{prefix}_{name}_ncu.txt")
        with open(path, "w") as f:
            f.write("\n".join(lines) + "\n")
        print("wrote", path)


if __name__ == "__main__":
    main()
