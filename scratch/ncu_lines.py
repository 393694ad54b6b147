import csv, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(['ncu','-i',rep,'--page','raw','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(raw.splitlines())); hdr,units,vals=rows[0],rows[1],rows[2]
keys=['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','smsp__issue_active.avg.pct_of_peak_sustained_active','sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','launch__grid_size','smsp__inst_executed.sum','sass__inst_executed_local_loads','sass__inst_executed_local_stores','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','launch__occupancy_limit_registers','launch__occupancy_limit_shared_mem','smsp__thread_inst_executed_per_inst_executed.ratio']
for h,u,v in zip(hdr,units,vals):
    if h in keys: print(h,u,v)
src = subprocess.run(['ncu','-i',rep,'--page','source','--print-source','cuda,sass','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(src.splitlines()))
cur=None; agg={}; H=None
for r in rows:
    if len(r)==2 and r[0]=='File Path': cur=r[1].split('/')[-1]; continue
    if len(r)>2 and r[0]=='Line No': H=r; continue
    if H is None or len(r)<10 or r[2]!='-': continue
    try: ln=int(r[0]); samples=int(r[4]); inst=int(r[7])
    except: continue
    if inst==0 and samples==0: continue
    st={}
    for name in ('stall_barrier','stall_long_sb','stall_short_sb','stall_wait','stall_math','stall_mio','stall_lg','stall_no_inst','stall_not_selected','stall_selected','stall_branch_resolving','stall_sleep','stall_membar','stall_dispatch'):
        if name in H:
            try: st[name]=int(r[H.index(name)])
            except: pass
    agg[(cur,ln)]=(samples,inst,r[1],st)
ts=sum(v[0] for v in agg.values()); ti=sum(v[1] for v in agg.values())
print('total samples',ts,'inst',ti)
tot={}
for v in agg.values():
    for k,x in v[3].items(): tot[k]=tot.get(k,0)+x
print('stalls:',{k:round(100*x/max(1,ts),1) for k,x in sorted(tot.items(),key=lambda kv:-kv[1])})
print('--- by samples')
for k,v in sorted(agg.items(), key=lambda kv:-kv[1][0])[:top]:
    big=max(v[3].items(),key=lambda kv:kv[1])[0] if v[3] else ''
    print(f"{k[0][:14]}:{k[1]:5d} samp {100*v[0]/ts:5.1f}% inst {100*v[1]/ti:5.1f}% {big[6:]:>10}  {v[2][:95]}")
