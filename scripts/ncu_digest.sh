#!/bin/bash
# usage: scripts/ncu_digest.sh gpurun_out/<tag>.ncu-rep [topN]  -> key metrics + top stalled SASS lines
rep=$1; topn=${2:-30}
ncu -i "$rep" --page raw --csv 2>/dev/null | python -c '
import csv,sys
rows=list(csv.reader(sys.stdin))
hdr=rows[0]; vals=rows[2] if len(rows)>2 else rows[1]
keep=["gpu__time_duration.sum","dram__bytes_read.sum","dram__bytes_write.sum","gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed","sm__warps_active.avg.pct_of_peak_sustained_active","launch__registers_per_thread","launch__occupancy_limit_registers","launch__occupancy_limit_shared_mem","smsp__issue_active.avg.pct_of_peak_sustained_active","smsp__inst_executed.sum","sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active","sm__inst_executed_pipe_adu.avg.pct_of_peak_sustained_active","l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed","l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum","l1tex__data_pipe_lsu_wavefronts_mem_shared.sum","smsp__thread_inst_executed_per_inst_executed.ratio","launch__grid_size","launch__block_size","lts__t_sectors_op_write.sum","lts__t_sectors_op_read.sum","l1tex__t_requests_pipe_lsu_mem_global_op_st.sum","l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum","l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum","l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum"]
for k in keep:
    if k in hdr: print(f"{k:75s} {vals[hdr.index(k)]} {rows[1][hdr.index(k)]}")
'
ncu -i "$rep" --page source --csv 2>/dev/null > /tmp/ncu_src.csv
python "$(dirname "$0")/ncu_top_stalls.py" /tmp/ncu_src.csv "$topn"
