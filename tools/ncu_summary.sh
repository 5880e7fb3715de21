#!/bin/bash
# tools/ncu_summary.sh gpurun_out/tag  -> key counters + per-function split of an ncu report
rep=$1
ncu -i $rep.ncu-rep --page raw --csv > ${rep}_raw.csv 2>/dev/null
python - <<PY
import csv
rows=list(csv.reader(open("${rep}_raw.csv")))
hdr=rows[0]; vals=rows[2]
want=['gpu__time_duration.sum','smsp__inst_executed.sum','smsp__issue_active.avg.pct_of_peak_sustained_active','sm__warps_active.avg.per_cycle_active','smsp__thread_inst_executed_per_inst_executed.ratio','launch__registers_per_thread','launch__block_size','sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','dram__bytes_read.sum','dram__bytes_write.sum']
for h,v in zip(hdr,vals):
    if h in want: print("%-70s %s" % (h, v))
st=[(float(v),h.split('issue_stalled_')[1].split('_per_issue')[0]) for h,v in zip(hdr,vals) if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('per_issue_active.ratio')]
print("stalls/issue: " + ", ".join("%s %.2f" % (n,x) for x,n in sorted(st,reverse=True)[:9]))
PY
ncu -i $rep.ncu-rep --page source --csv --print-source sass,cuda > ${rep}_src.csv 2>/dev/null
python tools/ncu_by_function.py ${rep}_src.csv --lines ${2:-25}
