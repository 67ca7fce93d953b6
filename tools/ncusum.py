#!/usr/bin/env python
"""usage: tools/ncusum.py <file.ncu-rep>  -- one-screen summary of the first captured kernel:
duration, DRAM bytes/throughput, issue utilisation, pipes, stall reasons per issue."""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h = rows[0]
for r in rows[2:]:
    d = dict(zip(h, r))
    print(d.get("Kernel Name"), "regs", d.get("launch__registers_per_thread"))
    keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "dram__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.sum", "smsp__inst_executed.sum",
            "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
            "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "sm__icc_request_hit_rate.pct",
            "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct"]
    for k in keys:
        if k in d:
            print(f"  {k:75s} {d[k]}")
    for k, v in d.items():
        if "issue_stalled" in k and k.endswith("per_issue_active.ratio") and "not_issued" not in k:
            try:
                if float(v) > 0.05:
                    print(f"  stall {k.split('issue_stalled_')[-1][:-len('_per_issue_active.ratio')]:28s} {float(v):.2f}")
            except ValueError:
                pass
