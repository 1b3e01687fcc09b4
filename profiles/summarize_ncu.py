import csv,sys,subprocess
rep=sys.argv[1]
raw=subprocess.run(['ncu','-i',rep,'--page','raw','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(raw.splitlines()))
hdr=rows[0]; units=rows[1]; r=rows[2]
want=['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','launch__occupancy_limit_registers','launch__occupancy_limit_shared_mem','lts__t_sector_hit_rate.pct','smsp__issue_active.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum','launch__grid_size','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio']
for w in want:
    for i,h in enumerate(hdr):
        if h==w: print(w, r[i], units[i])
src=subprocess.run(['ncu','-i',rep,'--page','source','--csv','--print-source','cuda,sass','--launch-count','1'],capture_output=True,text=True).stdout
rows=list(csv.reader(src.splitlines()))
hi=[i for i,r in enumerate(rows) if r and r[0]=='Line No'][0]
hdr=rows[hi]; ci=hdr.index('Instructions Executed'); si=hdr.index('# Samples')
lines=[];cur=None;tot=0;tots=0
for r in rows:
    if r and r[0]=='File Path': cur=r[1].split('/')[-1]
    if len(r)>ci and r[0] not in ('','Line No'):
        try: inst=int(r[ci]); s=int(r[si])
        except: continue
        lines.append((inst,s,cur,r[0],r[1][:100])); tot+=inst; tots+=s
print('total inst',tot,'samples',tots)
N=int(sys.argv[2]) if len(sys.argv)>2 else 25
for l in sorted(lines,reverse=True)[:N]: print('%5.1f%% inst %5.1f%% smp  %s:%s  %s'%(100*l[0]/tot,100*l[1]/tots,l[2],l[3],l[4]))
print('--- by samples')
for l in sorted(lines,key=lambda x:-x[1])[:12]: print('%5.1f%% inst %5.1f%% smp  %s:%s  %s'%(100*l[0]/tot,100*l[1]/tots,l[2],l[3],l[4]))
