import csv,sys
rows=list(csv.reader(open(sys.argv[1])))
h=rows[0]; v=rows[2]
want=['gpu__time_duration.sum','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum','l1tex__data_pipe_lsu_wavefronts','smsp__average_warps_issue_stalled','launch__registers_per_thread ','subpipe_dmma.avg','sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed','dram__bytes_read.sum ','dram__bytes_write.sum ','hit_rate','sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_elapsed','l1tex__t_requests_pipe_lsu_mem_global_op','l1tex__t_output_wavefronts_pipe_lsu_mem_global_op','bank_conflicts','smsp__inst_executed_op_shared','smsp__inst_executed_op_global','sm__inst_executed_pipe_lsu','smsp__cycles_active.avg ']
for name,val in zip(h,v):
    if any(w in name for w in want) and 'per_second' not in name and 'peak_sustained ' not in name and '.max' not in name and '.min' not in name:
        try:
            if float(val)==0: continue
        except: pass
        print(name, val)
