import csv,sys,subprocess
rep=sys.argv[1]
out=subprocess.run(['ncu','-i',rep,'--page','raw','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
hdr=rows[0]
for ri in range(2,len(rows)):
    val=rows[ri]
    d=dict(zip(hdr,val)); u=dict(zip(hdr,rows[1]))
    print('==', d.get('Kernel Name'), d.get('Grid Size'))
    keys=['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','sm__throughput.avg.pct_of_peak_sustained_elapsed','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__issue_active.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum','smsp__thread_inst_executed.sum','launch__registers_per_thread','launch__occupancy_limit_registers','launch__occupancy_limit_shared_mem','launch__waves_per_multiprocessor','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','lts__t_bytes.sum','lts__t_sectors_op_red.sum','lts__t_sectors_op_atom.sum','sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_adu.avg.pct_of_peak_sustained_active','smsp__average_warp_latency_per_inst_issued.ratio','smsp__cycles_active.avg','sm__cycles_elapsed.avg','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','smsp__inst_executed_op_shared_ld.sum','smsp__inst_executed_op_shared_st.sum']
    for k in keys:
        if k in d: print(f'  {k} [{u[k]}] = {d[k]}')
    st=[(float(d[k]),k) for k in hdr if k.startswith('smsp__average_warps_issue_stalled') and k.endswith('per_issue_active.ratio') and d[k] not in ('','n/a')]
    for v,k in sorted(st,reverse=True)[:10]: print(f'  stall {k[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]} = {v:.2f}')
