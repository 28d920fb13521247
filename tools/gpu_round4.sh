#!/bin/bash
# A/B: translated leaves with user functions inlined (default build) vs as real calls (libza_gpu_noinl.so)
mkdir -p gpurun_out; rm -f gpurun_out/ab_times.log
for k in RTT SOMA ATTACK ERBTilt SpectralStabilizer DDT DPT SaliencePush ADS BedRock Roomalizer TSEQ; do
  for v in inl noinl; do
    if [ $v = noinl ]; then export ZA_GPU_LIB=/root/repo/zorakaudio-experimental-plugins_b200/za_gpu/libza_gpu_noinl.so; else unset ZA_GPU_LIB; fi
    echo "== $k $v $(timeout 300 python tools/prof_run.py --plugin $k --streams 4096 --seconds 1 --reps 2 2>&1 | tail -1 | cut -c1-200)" >> gpurun_out/ab_times.log
  done
done
unset ZA_GPU_LIB
cat gpurun_out/ab_times.log
