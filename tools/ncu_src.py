#!/usr/bin/env python
"""Summarise an ncu source-page CSV: stall reasons and the hottest SASS lines."""
import csv, sys
f=sys.argv[1]; ntop=int(sys.argv[2]) if len(sys.argv)>2 else 25
rows=list(csv.reader(open(f)))
hdr=rows[1]; data=[r for r in rows[2:] if len(r)==len(hdr)]
si=hdr.index('# Samples'); src=hdr.index('Source')
stalls=[i for i,h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
def iv(x):
    try: return int(float(x))
    except: return 0
tot=sum(iv(r[si]) for r in data)
print(f,'total samples',tot, 'instructions', len(data))
agg={hdr[i]:sum(iv(r[i]) for r in data) for i in stalls}
print(sorted(agg.items(), key=lambda x:-x[1])[:8])
for r in sorted(data,key=lambda r:-iv(r[si]))[:ntop]:
    st=sorted([(iv(r[i]),hdr[i]) for i in stalls],reverse=True)[:2]
    print(r[si], r[src][:100], st)
