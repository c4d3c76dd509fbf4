import csv,sys,subprocess
rep=sys.argv[1]
out=subprocess.run(['ncu','-i',rep,'--page','raw','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
hdr=rows[0]
def col(n): return hdr.index(n) if n in hdr else None
want=[('gpu__time_duration.sum','us'),('dram__bytes_read.sum','rd'),('dram__bytes_write.sum','wr'),('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','dram%'),('sm__throughput.avg.pct_of_peak_sustained_elapsed','sm%'),('smsp__issue_active.avg.pct_of_peak_sustained_active','issue%'),('sm__warps_active.avg.pct_of_peak_sustained_active','occ%'),('launch__registers_per_thread','regs'),('l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','bankconf'),('smsp__inst_executed.sum','inst'),('lts__t_sector_hit_rate.pct','l2hit'),('l1tex__throughput.avg.pct_of_peak_sustained_active','l1%'),('sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active','lsu%'),('sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active','fma%'),('sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active','alu%'),('sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active','fp64%'),('sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active','xu%')]
units=rows[1]
for r in rows[2:]:
    name=r[col('Kernel Name')][:28]; grid=r[col('Grid Size')]
    vals=[]
    for m,lab in want:
        c=col(m)
        if c is None: continue
        v=r[c]
        try:
            f=float(v.replace(',',''))
            u=units[c]
            if lab in('rd','wr'):
                f=f*{'Kbyte':1e-3,'Mbyte':1,'Gbyte':1e3,'byte':1e-6}.get(u,1); v=f'{f:.1f}MB'
            elif lab=='us':
                f=f*{'us':1,'ms':1e3,'ns':1e-3}.get(u,1); v=f'{f:.1f}'
            elif lab in('inst','bankconf'): v=f'{f/1e6:.2f}M'
            else: v=f'{f:.1f}'
        except: pass
        vals.append(f'{lab}={v}')
    st=[(h.replace('smsp__average_warps_issue_stalled_','').replace('_per_issue_active.ratio',''),float(r[i])) for i,h in enumerate(hdr) if 'issue_stalled' in h and h.endswith('per_issue_active.ratio') and 'not_issued' not in h]
    st=sorted(st,key=lambda kv:-kv[1])[:5]
    print(name,grid,' '.join(vals)); print('     stalls:',' '.join(f'{k}={v:.2f}' for k,v in st))
