import csv, sys
path = sys.argv[1]; B = int(sys.argv[2]) if len(sys.argv) > 2 else 100
rows=list(csv.reader(open(path)))
H=rows[1]; ie=H.index('Instructions Executed'); te=H.index('Thread Instructions Executed'); so=H.index('Source'); sa=H.index('# Samples')
data=[(r[so],float(r[ie] or 0),float(r[te] or 0),float(r[sa] or 0)) for r in rows[2:] if len(r)>te]
tot=sum(d[1] for d in data); tots=sum(d[3] for d in data)
print(len(data),'instrs; total warp instr %.1f G; thread instr %.1f G'%(tot/1e9, sum(d[2] for d in data)/1e9))
for i in range(0,len(data),B):
    ch=data[i:i+B]; s=sum(d[1] for d in ch); t=sum(d[2] for d in ch); sm=sum(d[3] for d in ch)
    ops={}
    for d in ch:
        w=d[0].split()
        op=w[1] if w[0].startswith('@') else w[0]
        ops[op.split('.')[0]]=ops.get(op.split('.')[0],0)+d[1]
    top=sorted(ops.items(),key=lambda x:-x[1])[:6]
    print('%5d %5.1f%% instr %5.1f%% samples  avg thr %4.1f  '%(i,100*s/tot,100*sm/tots,t/max(s,1)), ' '.join('%s:%.0f%%'%(k,100*v/max(s,1)) for k,v in top))
