import csv, sys
rows=list(csv.reader(open(sys.argv[1])))
W=int(sys.argv[2]) if len(sys.argv)>2 else 40
# find header rows (there can be several kernels)
hi=[i for i,r in enumerate(rows) if 'Instructions Executed' in r]
for n,h in enumerate(hi):
    H=rows[h]; end=hi[n+1]-1 if n+1<len(hi) else len(rows)
    iS=H.index('Source'); iE=H.index('Instructions Executed'); iSamp=H.index('# Samples'); iT=H.index('Avg. Threads Executed')
    data=[]
    for r in rows[h+1:end]:
        try: data.append((r[iS].strip(), int(r[iE]), int(r[iSamp]), float(r[iT])))
        except Exception: pass
    tot=sum(d[1] for d in data); totS=sum(d[2] for d in data)
    print('kernel', rows[h-1][1][:90] if h>0 else '', len(data), tot, totS)
    for i in range(0,len(data),W):
        seg=data[i:i+W]
        e=sum(d[1] for d in seg); s=sum(d[2] for d in seg)
        if e/tot<0.004: continue
        tag=[k for k in ('VABSDIFF4','VIMNMX.S16x2','POPC','LDG','STG','SHFL','LDS','STS','VOTE','ATOMS','WARPSYNC','MUFU','BAR') if any(k in d[0] for d in seg)]
        print(f"{i:5d} exec={e/tot*100:5.1f}% samp={s/totS*100:5.1f}% thr={sum(d[3] for d in seg)/len(seg):4.1f} maxexec={max(d[1] for d in seg)} {tag}")
