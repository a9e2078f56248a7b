import csv,collections,sys
rows=list(csv.reader(l for l in open(sys.argv[1]) if not l.startswith("==")))
h=rows[0]; ik=h.index("Kernel Name"); iv=h.index("Metric Value")
t=collections.Counter(); n=collections.Counter()
for r in rows[1:]:
    if len(r)<=iv: continue
    k=r[ik].split("(")[0][:60]; t[k]+=float(r[iv].replace(",","")); n[k]+=1
for k,v in t.most_common(6): print("%-62s %4d %10.2f ms  %8.3f ms each"%(k,n[k],v/1e6,v/1e6/n[k]))
