#!/usr/bin/env python
"""Key counters of one kernel launch from an ncu --set full report, as text (for profiles/).

usage: ncu_summary.py <report.ncu-rep>
"""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__block_size", "launch__grid_size",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__shared_mem_per_block_static",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        # north star: L2 sector efficiency and FP64 pipe utilisation
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sectors.sum", "lts__t_sectors_srcunit_tex_op_read.sum",
        "lts__t_sectors_srcunit_tex_op_write.sum", "lts__t_sector_op_read_hit_rate.pct", "dram__sectors_read.sum", "dram__sectors_write.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum"]
STALL = "smsp__average_warps_issue_stalled_"


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, u = rows[0], rows[1]
    for v in rows[2:]:
        d = dict(zip(h, v))
        print("kernel:", d.get("Kernel Name"), " grid", d.get("Grid Size"), " block", d.get("Block Size"))
        for k in WANT:
            if k in d:
                print("  %-70s %-16s %s" % (k, u[h.index(k)], d[k]))
        stalls = sorted(((float(d[k]), k) for k in h if k.startswith(STALL) and k.endswith("_per_issue_active.ratio") and d[k]),
                        reverse=True)
        print("  warp stall reasons (warps stalled per issue-active cycle):")
        for x, k in stalls[:8]:
            print("    %-28s %.3f" % (k[len(STALL):-len("_per_issue_active.ratio")], x))


if __name__ == "__main__":
    main()
