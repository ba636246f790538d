import sys, time, os, ctypes as C
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(),'tests'))
import numpy as np
import support as S
import otter_b200 as ob
P = S.port()
rng = np.random.default_rng(3)
reads=[]; rows=[]
for tl,e,xf in [(100,0.02,0.2),(300,0.05,0.4),(300,0.15,0.5),(1500,0.005,0.2),(1500,0.05,0.3),(3000,0.05,0.2),(800,0.3,0.3),(8000,0.05,0.2)]:
    for reg in range(3):
        rr=S.tr_region(rng,5,tl,e,extra_frac=xf,n_rate=0.1)
        base=len(reads); reads+=rr
        for i in range(4):
            a,b=base+i,base+i+1
            d=len(reads[a])-len(reads[b])
            rows.append((a,b,-1,0,0,0))
            if d>=0: rows+=[(a,b,0,d,0,0),(a,b,d,0,0,0),(a,b,d//2,d//2,0,0)]
            else: rows+=[(a,b,0,0,0,-d),(a,b,0,0,-d,0)]
bases,off,lens=S.pack_reads(reads)
pairs=np.array(rows,dtype=ob.capi.PAIR_DTYPE)
ctx=ob.Context(0)
t0=time.time()
scores,cigs=ctx.affine_cigars(bases,off,lens,pairs)
print("gpu",time.time()-t0,ctx.stats())
bad=0
t0=time.time()
for i,(p,t,pb,pe,tb,te) in enumerate(pairs.tolist()):
    x,y=S.b(reads[p]),S.b(reads[t])
    buf=C.create_string_buffer(len(x)+len(y)+8); sc=C.c_int()
    n=P.wfo_affine_cigar(x,len(x),y,len(y),int(pb>=0),max(pb,0),pe,tb,te,4,6,2,C.byref(sc),buf,len(buf))
    exp=buf.raw[:n]
    if scores[i]!=sc.value or cigs[i]!=exp:
        bad+=1
        if bad<5: print("MISMATCH",i,(len(x),len(y),pb,pe,tb,te),scores[i],sc.value, (len(cigs[i]) if isinstance(cigs[i],bytes) else cigs[i]), n)
print(len(pairs),"pairs",bad,"mismatches; cpu",time.time()-t0)
