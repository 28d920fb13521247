import sys
sys.path[:0]=['/root/repo/zorakaudio-experimental-plugins_b200','/root/repo/oracle']
import numpy as np, za_gpu, za_oracle as zo
from za_gpu import workloads as wl
key=sys.argv[1]; ns=int(sys.argv[2]); nf=int(sys.argv[3])
ref=zo.RefPlugin(key)
x=wl.programme_bursts(ns,nf,n_ch=ref.n_in)
base=np.array([s["default"] for s in ref.sliders],np.float32); p=np.tile(base,(ns,1))
y,yd,v=ref.render(p,x,want_f64=True,want_vars=True)
bp=za_gpu.BatchProcessor(key,ns,precision="f64").prepare(p)
tap=np.zeros_like(yd); out=bp.process(x,tap64=tap)
d=np.abs(tap-yd)
print("per-stream max d:", d.max(axis=(1,2))[:8], "ref peak", np.abs(yd).max())
fr=np.argmax(d[0].max(axis=0)>1e-12); print("first bad frame stream0:", fr, tap[0,:,fr], yd[0,:,fr])
try:
    names=list(v)[:400]
    bad=[]
    for n in names:
        try: g=bp.read_state(n)
        except Exception as e:
            if 'heap' in str(e): print("ERR:",e); break
            continue
        if abs(g[0]-v[n][0])>1e-9*max(1,abs(v[n][0])): bad.append((n,g[0],v[n][0]))
    print("bad vars",len(bad),bad[:12])
except Exception as e: print("EXC",e)
