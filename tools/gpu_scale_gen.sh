#!/bin/bash
# how the translated leaves scale with the stream count (one warp per 32 streams)
mkdir -p gpurun_out; rm -f gpurun_out/scale_gen.log
for k in SpectralStabilizer ERBTilt RTT; do
  for n in 2048 4096 8192 16384 37888; do
    echo "== $k $n $(timeout 300 python tools/prof_run.py --plugin $k --streams $n --seconds 0.5 --reps 2 2>&1 | tail -1 | cut -c1-200)" >> gpurun_out/scale_gen.log
  done
done
cat gpurun_out/scale_gen.log
