import sys; sys.path.insert(0,'oracle'); sys.path.insert(0,'zorakaudio-experimental-plugins_b200')
import numpy as np, za_oracle as zo, za_gpu
from za_gpu import workloads as wl
ns,nf=48,48000*2+13
p,x=wl.cascade8_params(ns),wl.noise_two_sines(ns,nf)
y,yd=zo.best_oracle('cascade8').render(p,x,want_f64=True)
for prec in ('f64','f32'):
  for w in (1,2,4):
    bp=za_gpu.BatchProcessor('cascade8',ns,precision=prec); bp.set_option('warps_per_stream',w); bp.prepare(p)
    if prec=='f64':
        tap=np.zeros_like(yd); out=bp.process(x,tap64=tap)
        e=np.abs(tap-yd).max(axis=(1,2))/np.abs(yd).max(axis=(1,2))
        print(prec,w,'max|d|/peak: max %.2e median %.2e'%(e.max(),np.median(e)),'float-equal %.5f'%np.mean(out==y))
    else:
        out=bp.process(x); r=out.astype(np.float64)-yd
        d=np.array([20*np.log10(np.sqrt((r[s]**2).mean())) for s in range(ns)])
        print(prec,w,'resid dBFS rms: min %.1f med %.1f max %.1f'%(d.min(),np.median(d),d.max()))
    bp.close()
