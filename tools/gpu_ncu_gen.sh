#!/bin/bash
# full captures of two heap-using translated leaves (what bounds a mem[]-heavy @sample loop?)
mkdir -p gpurun_out
for k in SpectralStabilizer ERBTilt; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:jsfx_process_kernel -c 1 -f -o gpurun_out/r01_gen_$k \
    python tools/prof_run.py --plugin $k --streams 4096 --seconds 0.5 --reps 1 > gpurun_out/ncu_$k.log 2>&1
done
ls -la gpurun_out/r01_gen_*.ncu-rep
