#!/usr/bin/env python
"""Per-CUDA-source-line instruction counts and stall samples from an ncu report (needs -lineinfo)."""
import csv, subprocess, sys
rep=sys.argv[1]; ntop=int(sys.argv[2]) if len(sys.argv)>2 else 30
out=subprocess.run(['ncu','-i',rep,'--page','source','--print-source','cuda,sass','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
cur=None; hdr=None; res={}; fn=None
for r in rows:
    if r and r[0]=="File Path": cur=r[1].split('/')[-1]; continue
    if r and r[0]=="Function Name": fn=r[1][:40]; continue
    if r and r[0]=="Line No": hdr=r; continue
    if hdr and len(r)==len(hdr) and r[0]!="":
        k=(fn,cur,int(r[0]),r[1].strip())
        a=res.setdefault(k,[0,0,{}])
        a[0]+=int(r[hdr.index('# Samples')] or 0); a[1]+=int(r[hdr.index('Instructions Executed')] or 0)
        for i,h in enumerate(hdr):
            if h.startswith('stall_') and 'Not Issued' not in h and r[i] not in ('','0'): a[2][h]=a[2].get(h,0)+int(r[i])
ts=sum(v[0] for v in res.values()); ti=sum(v[1] for v in res.values())
print('samples',ts,'warp-instr',ti)
agg={}
for v in res.values():
    for h,c in v[2].items(): agg[h]=agg.get(h,0)+c
print(sorted(agg.items(),key=lambda x:-x[1])[:9])
for k,v in sorted(res.items(),key=lambda kv:-kv[1][0])[:ntop]:
    st=sorted(v[2].items(),key=lambda x:-x[1])[:2]
    print(f"{k[1][:16]:16s} L{k[2]:4d} samp={100*v[0]/ts:5.1f}% inst={100*v[1]/ti:5.1f}% {st} | {k[3][:80]}")
