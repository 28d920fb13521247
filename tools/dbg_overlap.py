import sys, time
sys.path[:0]=['/root/repo/zorakaudio-experimental-plugins_b200']
import torch, za_gpu
from za_gpu import workloads as wl
def mk(key, ns, nf):
    info=za_gpu.plugin_info(key); st=torch.cuda.Stream()
    bp=za_gpu.BatchProcessor(key, ns, 48000.0, 512, "f32", 0); bp.set_stream(st.cuda_stream); bp.prepare(wl.default_params(key, ns))
    x=(torch.rand((ns,info.n_in,nf),device="cuda")-0.5)*0.5; y=torch.empty((ns,info.n_out,nf),device="cuda")
    return bp,st,x,y
jobs={k:mk(k,4096,24000) for k in ("SpectralStabilizer","ERBTilt","RTT")}
torch.cuda.synchronize()
def run(keys, reps=2):
    for k in keys: jobs[k][0].process(jobs[k][2], out=jobs[k][3])
    torch.cuda.synchronize(); t0=time.perf_counter()
    for _ in range(reps):
        for k in keys:
            t1=time.perf_counter(); jobs[k][0].process(jobs[k][2], out=jobs[k][3]); dt=time.perf_counter()-t1
            if dt>0.01: print("   host blocked %.0f ms in process(%s)"%(dt*1e3,k))
    torch.cuda.synchronize(); return (time.perf_counter()-t0)/reps*1e3
for ks in (["SpectralStabilizer"],["ERBTilt"],["RTT"],["SpectralStabilizer","ERBTilt","RTT"]):
    print(ks, "%.1f ms per step"%run(ks))
