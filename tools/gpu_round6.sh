#!/bin/bash
# fp32 fast division in the translated leaves: residuals (printed by the test) and timing
mkdir -p gpurun_out; rm -f gpurun_out/r6_times.log
timeout 900 python -m pytest tests/test_generated_gpu.py -m gpu -q -s -k "fp32" 2>&1 | grep -E "FP32-RESIDUAL|passed|failed|Error" > gpurun_out/tests6.log
for k in RTT SOMA ATTACK ERBTilt SpectralStabilizer DDT DPT SaliencePush ADS BedRock Roomalizer; do
  echo "== $k $(timeout 300 python tools/prof_run.py --plugin $k --streams 4096 --seconds 1 --reps 2 2>&1 | tail -1 | cut -c1-200)" >> gpurun_out/r6_times.log
done
cat gpurun_out/tests6.log gpurun_out/r6_times.log
