#!/bin/bash
# translated leaves after the register-resident context change: parity, then timing of every leaf
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_generated_gpu.py -m gpu -q -x 2>&1 | tail -4 > gpurun_out/tests3.log
for k in RTT SOMA ATTACK ERBTilt SpectralStabilizer DDT DPT SaliencePush ADS BedRock Roomalizer TSEQ; do
  echo "== $k" >> gpurun_out/gen_times.log
  timeout 300 python tools/prof_run.py --plugin $k --streams 4096 --seconds 1 --reps 2 2>&1 | tail -2 >> gpurun_out/gen_times.log
done
cat gpurun_out/tests3.log gpurun_out/gen_times.log
