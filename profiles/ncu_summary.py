import csv,sys,subprocess
want=['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','launch__grid_size','launch__block_size','launch__shared_mem_per_block_dynamic','launch__occupancy_limit_shared_mem','smsp__inst_executed.sum','smsp__issue_active.avg.pct_of_peak_sustained_active','smsp__thread_inst_executed_per_inst_executed.ratio','l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_st.sum','l1tex__t_requests_pipe_lsu_mem_global_op_st.sum','l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum','lts__throughput.avg.pct_of_peak_sustained_elapsed','sm__throughput.avg.pct_of_peak_sustained_elapsed']
stalls='smsp__average_warps_issue_stalled_%s_per_issue_active.ratio'
for s in ['short_scoreboard','long_scoreboard','wait','barrier','branch_resolving','not_selected','math_pipe_throttle','mio_throttle','lg_throttle','no_instruction','dispatch_stall']: want.append(stalls%s)
for f in sys.argv[1:]:
    out=subprocess.run(['ncu','-i',f,'--page','raw','--csv'],capture_output=True,text=True).stdout
    rows=list(csv.reader(out.splitlines()))
    hdr,units=rows[0],rows[1]
    for vals in rows[2:]:
        print('## %s  [%s]'%(vals[hdr.index('Kernel Name')],f.split('/')[-1]))
        for w in want:
            if w in hdr:
                i=hdr.index(w); print('%-90s %s %s'%(w,vals[i],units[i]))
        print()
