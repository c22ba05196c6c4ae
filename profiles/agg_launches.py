import csv, collections, re, os, sys
p=sys.argv[1]; nvc=float(sys.argv[2]) if len(sys.argv)>2 else 16
with open(p) as f:
    lines=[l for l in f if not l.startswith("==")]
allr=list(csv.DictReader(lines))
agg=collections.defaultdict(lambda:[0,0.0,1e9,0]); tot=0
for row in allr:
    name=row["Kernel Name"]; grid=row.get("Grid Size","")
    v=float(row["Metric Value"].replace(",","")); unit=row["Metric Unit"]
    v = v/1e6 if unit in("ns","nsecond") else v/1e3 if unit in ("us","usecond") else v
    if name.startswith("void at::") or "elementwise" in name: continue
    key=(re.sub(r"\(.*","",name)[:40], grid)
    agg[key][0]+=1; agg[key][1]+=v; tot+=v; agg[key][2]=min(agg[key][2],v); agg[key][3]=max(agg[key][3],v)
print("total ms per V-cycle",tot/nvc, "n", len(allr))
for k,v in sorted(agg.items(), key=lambda kv:-kv[1][1])[:int(sys.argv[3]) if len(sys.argv)>3 else 30]:
    print(f"{v[1]/nvc:8.3f} ms/vc {v[0]:5d} x  avg {v[1]/v[0]:8.4f} min {v[2]:8.4f} max {v[3]:8.4f} {k}")
