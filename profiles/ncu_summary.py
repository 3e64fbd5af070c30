import csv,collections,sys
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>10]
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value')
d=collections.defaultdict(list)
for r in rows[1:]:
    try: d[r[ki][:70]].append(float(r[vi].replace(',','')))
    except: pass
for k,v in d.items(): print(f"{k:70s} n={len(v):4d} mean_ns={sum(v)/len(v):10.1f} min={min(v):10.1f} max={max(v):10.1f}")
