import csv,sys,subprocess
rep=sys.argv[1]
src=subprocess.run(['ncu','-i',rep,'--page','source','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(src.splitlines()))
hdr=rows[1]; ix={h:i for i,h in enumerate(hdr)}
data=rows[2:]
tot=sum(int(r[ix['# Samples']] or 0) for r in data)
mx=max(int(r[ix['Instructions Executed']] or 0) for r in data)
print(rows[0][1][:90]); print('total samples',tot,'loop instrs',sum(1 for r in data if int(r[ix['Instructions Executed']] or 0)>=mx*0.7))
stalls=[h for h in hdr if h.startswith('stall_') and 'Not' not in h]
agg={h[6:]:sum(int(r[ix[h]] or 0) for r in data) for h in stalls}
print({k:v for k,v in sorted(agg.items(),key=lambda x:-x[1]) if v})
top=sorted(data,key=lambda r:-int(r[ix['# Samples']] or 0))[:int(sys.argv[2]) if len(sys.argv)>2 else 12]
for r in top:
    print(r[ix['Address']][-5:], r[ix['# Samples']].rjust(5), {h[6:]:r[ix[h]] for h in stalls if r[ix[h]] not in ('0','')}, r[ix['Instructions Executed']].rjust(7), r[ix['Source']][:80])
raw=subprocess.run(['ncu','-i',rep,'--page','raw','--csv'],capture_output=True,text=True).stdout
rr=list(csv.reader(raw.splitlines())); d=dict(zip(rr[0],rr[2]))
for k in ['gpu__time_duration.sum','smsp__issue_active.avg.pct_of_peak_sustained_active','sm__warps_active.avg.pct_of_peak_sustained_active','dram__throughput.avg.pct_of_peak_sustained_elapsed','l1tex__t_sector_hit_rate.pct','lts__t_sector_hit_rate.pct','launch__registers_per_thread','launch__occupancy_limit_shared_mem','launch__occupancy_limit_registers','launch__grid_size','launch__waves_per_multiprocessor','smsp__inst_executed.sum','l1tex__m_xbar2l1tex_read_bytes.sum','dram__bytes_read.sum']:
    print(k, d.get(k))
