import csv, sys, collections, subprocess
rep=sys.argv[1]
src=subprocess.run(['ncu','-i',rep,'--page','source','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(src.splitlines()))
hdr=rows[1]; data=rows[2:]
ix={h:i for i,h in enumerate(hdr)}
stalls=[h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
tot=sum(int(r[ix['# Samples']]) for r in data)
print('total samples',tot, 'ninstr', len(data))
agg={s:sum(int(r[ix[s]]) for r in data) for s in stalls}
for s,v in sorted(agg.items(), key=lambda kv:-kv[1])[:8]: print(' ',s, v, f"{100*v/tot:.1f}%")
ops=collections.Counter()
for r in data:
    t=r[ix['Source']].strip().split()
    op=t[1] if t[0].startswith('@') else t[0]
    ops[op.split('.')[0]]+=int(r[ix['Instructions Executed']])
te=sum(ops.values())
NV=float(sys.argv[2]) if len(sys.argv)>2 else 168.69e6
print('total exec per vis',te/NV)
print(' '.join(f"{o}:{v/NV:.2f}" for o,v in ops.most_common(24)))
for n,r in enumerate(data): r.append(n)
top=sorted(data, key=lambda r:-int(r[ix['# Samples']]))[:int(sys.argv[3]) if len(sys.argv)>3 else 12]
for r in sorted(top, key=lambda r:r[-1]):
    st=sorted(((int(r[ix[s]]),s) for s in stalls), reverse=True)[:2]
    print(f"{r[-1]:5d} {int(r[ix['# Samples']]):7d} {int(r[ix['Instructions Executed']]):11d}  {r[ix['Source']].strip()[:56]:56s} {st}")
raw=subprocess.run(['ncu','-i',rep,'--page','raw','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(raw.splitlines()))
for h,u,v in zip(rows[0],rows[1],rows[2]):
    if h in ('gpu__time_duration.sum','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum','l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum','l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed','sm__inst_executed.avg.per_cycle_elapsed','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','dram__bytes_read.sum','dram__bytes_write.sum','sm__warps_active.avg.pct_of_peak_sustained_active','launch__grid_size','launch__registers_per_thread','sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active','sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active','sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active'):
        print(h,u,v)

# optional 4th argument: write the DRAM traffic of this launch for bench.py's roofline.traffic
if len(sys.argv) > 4:
    import json
    vals = {h: (u, v) for h, u, v in zip(rows[0], rows[1], rows[2])}
    def to_bytes(name):
        u, v = vals[name]
        return float(v.replace(',', '')) * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[u]
    d = {'kernel': vals.get('Kernel Name', ('', ''))[1], 'dram_bytes': int(to_bytes('dram__bytes_read.sum') + to_bytes('dram__bytes_write.sum')),
         'source': rep}
    open(sys.argv[4], 'w').write(json.dumps(d) + '\n')
