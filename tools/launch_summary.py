import csv, collections, re, sys
with open(sys.argv[1]) as f:
    lines=[l for l in f if not l.startswith('==')]
r=list(csv.DictReader(lines))
agg=collections.defaultdict(lambda:[0,0.0])
for x in r:
    if x['Metric Name']!='gpu__time_duration.sum': continue
    k=re.sub(r'\(.*','',x['Kernel Name'])
    v=float(x['Metric Value'].replace(',',''))
    u=x['Metric Unit']
    v = v/1e3 if u in('ns','nsecond') else (v if u in ('us','usecond') else v*1e3)
    agg[k][0]+=1; agg[k][1]+=v
tot=sum(v[1] for v in agg.values())
print("total us",tot, "launches", sum(v[0] for v in agg.values()))
for k,v in sorted(agg.items(), key=lambda kv:-kv[1][1])[:int(sys.argv[2]) if len(sys.argv)>2 else 20]:
    print(f"{k[:60]:60s} n={v[0]:5d} us={v[1]:10.1f} share={v[1]/tot:.3f} avg={v[1]/v[0]:.1f}")
