#!/bin/bash
# what saturates when a heap leaf's single-warp CTAs are packed 3.5 per SM? (SpectralStabilizer, 16 384 streams)
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:jsfx_process_kernel -c 1 -f -o gpurun_out/r01_gen_ss_16k \
  python tools/prof_run.py --plugin SpectralStabilizer --streams 16384 --seconds 0.25 --reps 1 > gpurun_out/ncu_ss16k.log 2>&1
ls -la gpurun_out/r01_gen_ss_16k.ncu-rep
