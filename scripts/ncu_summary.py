#!/usr/bin/env python
"""Condense an `ncu --set full` report into the per-launch table kept under profiles/.
usage: python scripts/ncu_summary.py <report.ncu-rep> > profiles/rNN_<tag>.txt"""
import csv, io, subprocess, sys

KEEP = ["dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "sm__cycles_elapsed.avg", "smsp__inst_executed.sum",
        "sm__inst_issued.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tmem.avg.pct_of_peak_sustained_active", "smsp__sass_inst_executed_op_shared_ld.sum",
        "smsp__sass_inst_executed_op_tmem_ldt.sum", "smsp__sass_inst_executed_op_tmem_stt.sum",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active"]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units, data = rows[0], rows[1], rows[2:]


def line(name, unit, vals):
    print("%-84s %-8s %s" % (name, unit, "  ".join(vals)))


line("Kernel Name", "", [r[hdr.index("Kernel Name")][:14] for r in data])
for k in KEEP:
    if k in hdr:
        i = hdr.index(k)
        line(k, units[i], [r[i] for r in data])
pre, suf = "smsp__average_warps_issue_stalled_", "_per_issue_active.ratio"
for i, h in enumerate(hdr):
    if h.startswith(pre) and h.endswith(suf):
        line("stall:" + h[len(pre):-len(suf)], units[i], ["%.6f" % float(r[i].replace(",", "")) for r in data])
