"""Prints the judged metrics of an .ncu-rep (ncu --set full) capture: python profiles/ncu_summary.py <file.ncu-rep>"""
import csv,sys,subprocess
rep=sys.argv[1]
out=subprocess.run(['ncu','-i',rep,'--page','raw','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
hdr=rows[0]; units=rows[1]
keys=['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','launch__grid_size','launch__block_size','smsp__inst_executed.sum','smsp__issue_active.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active','smsp__warps_eligible.avg.per_cycle_active','l1tex__t_bytes_pipe_lsu_mem_global_op_ld.sum','l1tex__t_bytes_pipe_lsu_mem_global_op_st.sum','lts__t_bytes.sum','lts__t_sector_hit_rate.pct','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','smsp__inst_executed_op_shared_ld.sum','sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active','l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_active', 'lts__t_sectors_srcunit_tex_op_write.sum','lts__t_sectors_srcunit_tex_op_read.sum']
for r in rows[2:]:
    d=dict(zip(hdr,r))
    print('----',d.get('Kernel Name','')[:90])
    for k in keys:
        if k in d: print('  ',k, d[k], units[hdr.index(k)])
    for k in hdr:
        if k.startswith('smsp__average_warps_issue_stalled') and k.endswith('per_issue_active.ratio'):
            try:
                if float(d[k])>0.3: print('     stall',k.replace('smsp__average_warps_issue_stalled_','').replace('_per_issue_active.ratio',''),d[k])
            except: pass
