import csv, sys, collections
rows=list(csv.reader(open(sys.argv[1])))
top=int(sys.argv[2]) if len(sys.argv)>2 else 40
cur_file=None; agg=collections.OrderedDict(); hdr=None; kernel=None; first_kernel=None
for r in rows:
    if not r: continue
    if r[0]=='File Path': cur_file=r[1].split('/')[-1]; continue
    if r[0]=='Function Name':
        kernel=r[1]; 
        if first_kernel is None: first_kernel=kernel
        continue
    if r[0]=='Line No': hdr=r; ie=hdr.index('Instructions Executed'); sm=hdr.index('# Samples'); continue
    if hdr is None or kernel!=first_kernel: continue
    # cuda line rows have Line No filled; sass rows have Address
    try:
        if r[0]!='':
            cur_line=(cur_file,int(r[0]),r[1]); agg.setdefault(cur_line,[0,0])
        if len(r)>ie and r[2]!='' :
            agg[cur_line][0]+=int(r[ie] or 0); agg[cur_line][1]+=int(r[sm] or 0)
    except Exception as e: pass
tot=sum(v[0] for v in agg.values()); ts=sum(v[1] for v in agg.values())
print(first_kernel[:60], "total warp-inst", tot, "samples", ts)
for k,v in sorted(agg.items(), key=lambda kv:-kv[1][0])[:top]:
    print(f"{100*v[0]/tot:5.1f}%i {100*v[1]/max(ts,1):5.1f}%s {k[0]}:{k[1]:<4} {k[2].strip()[:100]}")
