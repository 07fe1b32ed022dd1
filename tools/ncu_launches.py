import csv, collections, re, sys
lines=[l for l in open(sys.argv[1]) if not l.startswith('==')]
r=csv.DictReader(lines)
agg=collections.defaultdict(lambda:[0,0.0])
for row in r:
    if row.get('Metric Name')!='gpu__time_duration.sum': continue
    k=row['Kernel Name']
    k=re.sub(r'void |cip::<unnamed>::|\(anonymous namespace\)::','',k)
    k=re.sub(r'\(.*','',k)[:80]
    v=float(row['Metric Value'].replace(',',''))
    u=row['Metric Unit']
    if u=='ns': v/=1e6
    elif u=='us': v/=1e3
    agg[k][0]+=1; agg[k][1]+=v
tot=sum(v[1] for v in agg.values())
for k,v in sorted(agg.items(), key=lambda kv:-kv[1][1])[:30]:
    print(f"{v[1]:10.3f} ms {v[0]:4d}  {100*v[1]/tot:5.1f}%  {k}")
print(tot)
