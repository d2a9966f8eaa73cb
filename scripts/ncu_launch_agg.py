import csv,collections,sys
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>5]
hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value"); ui=hdr.index("Metric Unit")
agg=collections.defaultdict(lambda:[0,0.0])
for r in rows[1:]:
    v=float(r[vi].replace(",","")); u=r[ui]
    v = v/1e3 if u in ("ns","nsecond") else (v*1e3 if u in ("ms","msecond") else v)
    agg[r[ki][:70]][0]+=1; agg[r[ki][:70]][1]+=v
tot=sum(t for n,t in agg.values())
print(f"total {tot/1e3:.3f} ms")
for k,(n,t) in sorted(agg.items(), key=lambda kv:-kv[1][1])[:int(sys.argv[2]) if len(sys.argv)>2 else 16]: print(f"{t/1e3:9.3f} ms {n:6d}  {k}")
