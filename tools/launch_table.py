import csv, sys
from collections import defaultdict
def table(path, div):
    rows=list(csv.reader(l for l in open(path) if l.startswith('"')))
    hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value")
    d=defaultdict(list)
    for r in rows[1:]:
        d[r[ki][:64]].append(float(r[vi].replace(',','')))
    tot=sum(sum(v) for v in d.values())
    for k,v in sorted(d.items(), key=lambda kv:-sum(kv[1])): print("%-66s n/step=%6.1f us/step=%9.1f avg=%8.1f  %5.1f%%"%(k,len(v)/div,sum(v)/1e3/div,sum(v)/len(v)/1e3,100*sum(v)/tot))
    print("total us/step", tot/1e3/div)
table(sys.argv[1], float(sys.argv[2]))
