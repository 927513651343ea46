#!/bin/bash
# per-kernel summary CSV of an ncu report: tools/ncu_summary.sh report.ncu-rep > profiles/xxx.csv
ncu -i "$1" --page raw --csv --metrics gpu__time_duration.sum,launch__grid_size,launch__registers_per_thread,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,lts__throughput.avg.pct_of_peak_sustained_elapsed,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed 2>/dev/null | python3 -c "
import csv,sys
r=list(csv.reader(sys.stdin))
h=r[0]; keep=[h.index('Kernel Name')]+[i for i,c in enumerate(h) if '__' in c]
for row in r:
    print(','.join(row[i].split('(')[0].replace('orbx::','') if i==keep[0] else row[i] for i in keep))
"
