#!/usr/bin/env python
"""Key figures of every launch in an .ncu-rep (`ncu --set full`): duration, DRAM bytes, throughputs, instruction mix,
stall-sample shares.  python tools/ncu_one.py file.ncu-rep ["title line"]"""
import csv
import io
import subprocess
import sys

KEYS = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_active", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum",
        "smsp__sass_inst_executed_op_shared.sum", "smsp__sass_inst_executed_op_global_ld.sum",
        "l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum"]


def main(path, title=None):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    if title:
        print(title)
    for r in rows[2:]:
        for k in KEYS:
            if k in idx:
                print(f"   {k:70s} {r[idx[k]]:>22s} {units[idx[k]]}")
        st = [(float(r[i].replace(",", "") or 0), h) for h, i in idx.items()
              if h.startswith("smsp__pcsamp_warps_issue_stalled") and not h.endswith("not_issued")]
        tot = sum(v for v, _ in st) or 1.0
        print("   stall samples (share):")
        for v, h in sorted(st, reverse=True)[:8]:
            print(f"      {h[len('smsp__pcsamp_warps_issue_stalled_'):]:28s} {100 * v / tot:5.1f} %")


if __name__ == "__main__":
    main(*sys.argv[1:3])
