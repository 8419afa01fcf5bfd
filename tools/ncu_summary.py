"""Condense `ncu --page raw --csv` output to the columns DESIGN/bench cite.

    python tools/ncu_summary.py raw.csv > profiles/rNN_ncu_full_summary.csv
"""
import csv
import sys

WANT = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "smsp__inst_executed.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem"]


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    idx = [hdr.index(w) for w in WANT if w in hdr]
    out = csv.writer(sys.stdout)
    out.writerow([hdr[i] for i in idx])
    out.writerow([units[i] for i in idx])
    for r in rows[2:]:
        v = [r[i] for i in idx]
        v[0] = v[0].split("(")[0]
        out.writerow(v)


if __name__ == "__main__":
    main()
