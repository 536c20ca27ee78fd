import csv, collections, sys
path=sys.argv[1]
lines=[l for l in open(path) if not l.startswith('==')]
rows=list(csv.DictReader(lines))
agg=collections.defaultdict(lambda:[0,0.0])
for row in rows:
    name=row['Kernel Name']; v=float(row['Metric Value'].replace(',','')); unit=row['Metric Unit']
    if unit=='ns': v/=1e3
    elif unit=='ms': v*=1e3
    agg[name][0]+=1; agg[name][1]+=v
tot=sum(v[1] for v in agg.values())
nsteps=float(sys.argv[2]) if len(sys.argv)>2 else 7
print('total us',tot,'per step ms',tot/nsteps/1e3)
for k,v in sorted(agg.items(), key=lambda kv:-kv[1][1])[:40]:
    print('%-64s n/step=%5.1f  ms/step=%7.3f  %5.1f%%  avg_us=%8.1f'%(k[:64],v[0]/nsteps,v[1]/nsteps/1e3,100*v[1]/tot,v[1]/v[0]))
