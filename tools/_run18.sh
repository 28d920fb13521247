TAG=r02B PYTEST_ARGS="--maxfail=5" tools/gpu_run.sh tests
python - <<PY
import sys, time, numpy as np
sys.path.insert(0,'zorakaudio-experimental-plugins_b200'); sys.path.insert(0,'oracle')
import za_gpu, torch
from za_gpu import workloads as wl
import za_oracle as zo
ns, sec, block = 1024, 10, 512
nf = 48000*sec; nb = (nf+block-1)//block
ref = zo.best_oracle("cascade8")
p = wl.automation(ref.sliders, nb, ns, which=[0,4,14], base=wl.cascade8_params(ns)[0])
x = torch.rand((ns,1,nf), device="cuda")*0.2-0.1
for prec in ("f32","f64"):
    bp = za_gpu.BatchProcessor("cascade8", ns, 48000.0, block, prec).prepare(p[0])
    y = bp.process(x, params=p); bp.sync()
    t=time.time(); y = bp.process(x, params=p); bp.sync(); dt=time.time()-t
    ms, nl = bp.last_kernel_ms()
    print(f"automated cascade8 {prec}: {ns} streams x {sec} s, wall {dt*1e3:.1f} ms, kernels {ms:.1f} ms in {nl} launches -> {ns*nf/dt:.3g} frames/s")
    bp.close()
PY
