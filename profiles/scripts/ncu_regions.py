import csv, sys
rows=list(csv.reader(open(sys.argv[1])))
which=int(sys.argv[2]); nchunks=int(sys.argv[3])
ks=[]; cur=None
for r in rows:
    if r and r[0]=="Kernel Name": cur={'name':r[1],'rows':[]}; ks.append(cur); continue
    if r and r[0]=="Address": cur['hdr']=r; continue
    if cur is not None and r: cur['rows'].append(r)
kk=[x for x in ks if x["rows"]]; k=kk[min(which*2,len(kk)-1)] if len(kk)>2 else kk[which]
h={n:i for i,n in enumerate(k['hdr'])}
tot=sum(int(r[h['Instructions Executed']]) for r in k['rows'])
print(k['name'][:60], 'static', len(k['rows']), 'warp instr per chunk', tot/nchunks)
segs=[]; run_start=0; prev=None
for i,r in enumerate(k['rows']):
    e=int(r[h['Instructions Executed']])
    if prev is None or abs(e-prev)>0.03*max(e,prev,1):
        if prev is not None: segs.append((run_start,i-1,prev))
        run_start=i
    prev=e
segs.append((run_start,len(k['rows'])-1,prev))
for s,e,c in segs:
    w=(e-s+1)*c
    if w>0.012*tot:
        ops={}
        for r in k['rows'][s:e+1]:
            t=r[h['Source']].split()
            op=t[1] if t[0].startswith('@') else t[0]
            op=op.split('.')[0]
            ops[op]=ops.get(op,0)+1
        top=sorted(ops.items(),key=lambda x:-x[1])[:7]
        print('rows %4d-%4d n=%4d warps/chunk=%.2f share=%.1f%%'%(s,e,e-s+1,c/nchunks,100*w/tot), top)
