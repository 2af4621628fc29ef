"""Summarise an .ncu-rep (raw page) into the few metrics the roofline discussion needs.
usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep [pattern ...] > profiles/xxx.txt"""
import csv
import io
import subprocess
import sys

DEFAULT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_fp64", "sm__pipe_fp64_cycles_active",
           "sm__pipe_fmaheavy", "sm__inst_executed_pipe_fma", "launch__registers_per_thread", "launch__occupancy_limit",
           "sm__warps_active.avg.pct_of_peak_sustained_active",
           "l1tex__data_bank_conflicts_pipe_lsu_mem_shared", "smsp__pcsamp_warps_issue_stalled", "lts__t_sector_hit_rate.pct", "sm__cycles_elapsed.max",
           "smsp__inst_executed.sum", "launch__grid_size", "launch__block_size", "smsp__cycles_active.avg", "l1tex__t_bytes_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_bytes_pipe_lsu_mem_global_op_st.sum",
           "smsp__issue_active.avg.pct", "sm__pipe_tensor", "lts__t_bytes.sum ", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum",
           "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum"]


def main():
    rep = sys.argv[1]
    pats = sys.argv[2:] or DEFAULT
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ki = hdr.index("Kernel Name")
    print("kernels:", [r[ki][:60] for r in data])
    for i, h in enumerate(hdr):
        if any(p in h for p in pats):
            print(f"{h} [{units[i]}]: " + " | ".join(r[i] for r in data))


if __name__ == "__main__":
    main()
