#!/usr/bin/env python
"""Summarise one kernel of an .ncu-rep into the CSV format kept under profiles/:
python scripts/ncu_summary.py report.ncu-rep "header comment" > profiles/xyz_ncu_summary.csv"""
import csv
import subprocess
import sys

KEEP = """gpu__time_duration.sum launch__grid_size launch__block_size launch__registers_per_thread
launch__shared_mem_per_block_dynamic sm__warps_active.avg.pct_of_peak_sustained_active dram__bytes_read.sum
dram__bytes_write.sum lts__t_sector_hit_rate.pct lts__t_bytes.sum sm__throughput.avg.pct_of_peak_sustained_elapsed
smsp__issue_active.avg.pct_of_peak_sustained_active smsp__inst_executed.sum
smsp__thread_inst_executed_per_inst_executed.ratio sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active
sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active
l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed
l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed
l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum smsp__warps_eligible.avg.per_cycle_active""".split()

rep, note = sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else ""
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h, units, v = rows[0], rows[1], rows[2]
w = csv.writer(sys.stdout)
print(f'"# {note}"')
w.writerow(["kernel", v[h.index("Kernel Name")]])
w.writerow(["metric", "unit", "value"])
for k, u, x in zip(h, units, v):
    if k in KEEP or ("issue_stalled" in k and k.endswith("_per_issue_active.ratio")):
        w.writerow([k, u, x])
