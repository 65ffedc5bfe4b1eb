import csv, collections, sys, subprocess, io
rep, key = sys.argv[1], sys.argv[2]
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 30
out = subprocess.run(["ncu","-i",rep,"--page","source","--print-source","sass,cuda","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(io.StringIO(out)))
blocks=[]; cur=None; fname=None
for r in rows:
    if not r: continue
    if r[0] in ("File Path","File Name"): fname=r[1].split('/')[-1]; continue
    if r[0]=="Function Name":
        cur={'name':r[1],'hdr':None,'rows':[]}; blocks.append(cur); continue
    if cur is None: continue
    if r[0]=="Line No": cur['hdr']=r; continue
    cur['rows'].append((fname,r))
import os
SRC={f: open('/root/repo/do-you-really-need-to-pay-2-20-hedge-fund-strategy-replication-via-machine-learning_b200/csrc/'+f).read().split('\n') for f in ('fit_kernel.cu','fit_cluster.cu','common.cuh','fit_params.cuh','wide_kernel.cu','gen_tc.cu','tc.cuh','gen_kernel.cu')}
agg=collections.Counter(); inst=collections.Counter(); sass={}
for b in blocks:
    if key in b['name']:
        hdr=b['hdr']; si=hdr.index('# Samples'); ie=hdr.index('Instructions Executed'); ai=hdr.index('Address'); so=[i for i,h in enumerate(hdr) if h=='Source'][1]
        for f,r in b['rows']:
            if len(r)>si and r[si].isdigit():
                agg[(f,r[0])]+=int(r[si]); inst[(f,r[0])]+=int(r[ie]) if r[ie].isdigit() else 0
                a=r[ai]
                if a not in sass: sass[a]=[r[so], int(r[si]), int(r[ie]) if r[ie].isdigit() else 0, f, r[0]]
tot=sum(agg.values()); ti=sum(inst.values()); print(key, 'total samples', tot, 'total warp-inst', ti)
for (f,l),v in agg.most_common(topn):
    text = SRC[f][int(l)-1].strip()[:100] if f in SRC and l.isdigit() else ''
    print(f"  {v*100/tot:5.1f}%  inst={inst[(f,l)]*100/ti:5.1f}% {f}:{l:>4s}  {text}")
print("top sass")
for a,v in sorted(sass.items(), key=lambda kv:-kv[1][1])[:15]:
    print(f"  {v[1]*100/tot:5.2f}% inst={v[2]:>9d} {v[3]}:{v[4]:>4s} {v[0][:90]}")
print("opcode mix (warp instructions executed)")
ops=collections.Counter()
for a,v in sass.items():
    t=v[0].strip().split()
    if not t: continue
    op = t[1] if t[0].startswith('@') and len(t) > 1 else t[0]
    ops[op.split('.')[0]] += v[2]
for op,n in ops.most_common(28):
    print(f"  {n*100/ti:5.1f}%  {op}")
