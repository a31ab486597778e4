#!/usr/bin/env python
"""Hot SASS instructions of a kernel in an .ncu-rep (stall samples per instruction) + a few headline metrics.
usage: scripts/ncu_hot.py <report.ncu-rep> [min_fraction]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; frac = float(sys.argv[2]) if len(sys.argv) > 2 else 0.01
out = subprocess.run(f"ncu -i {rep} --page source --csv --print-source sass", shell=True, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]; ci = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[2:] if len(r) >= len(hdr)]
tot = sum(int(r[ci['# Samples']] or 0) for r in data)
print(rows[0][1][:80], 'instrs', len(data), 'samples', tot)
agg = {}
for k, r in enumerate(data):
    s = int(r[ci['# Samples']] or 0)
    for h in hdr:
        if h.startswith('stall_') and 'Not' not in h:
            agg[h] = agg.get(h, 0) + int(r[ci[h]] or 0)
    if s >= tot * frac:
        st = {h: int(r[ci[h]] or 0) for h in hdr if h.startswith('stall_') and 'Not' not in h}
        top = sorted(st.items(), key=lambda x: -x[1])[:2]
        print(k, r[ci['Source']][:64].ljust(64), s, r[ci['Instructions Executed']], top)
print(sorted(agg.items(), key=lambda x: -x[1])[:8])
out = subprocess.run(f"ncu -i {rep} --page raw --csv", shell=True, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out))); hdr = rows[0]; r = rows[2]
for k in ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed',
          'l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum',
          'l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
          'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'lts__t_sector_hit_rate.pct']:
    if k in hdr: print(k, r[hdr.index(k)])
