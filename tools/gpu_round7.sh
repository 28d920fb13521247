#!/bin/bash
# branch-free heap bounds + base-less addressing + fast fp32 division in the translated leaves: parity, bounce, timing
mkdir -p gpurun_out; rm -f gpurun_out/r7_times.log
timeout 1200 python -m pytest tests/test_generated_gpu.py tests/test_bounce.py -m gpu -q 2>&1 | tail -4 > gpurun_out/tests7.log
for k in RTT SOMA ATTACK ERBTilt SpectralStabilizer DDT DPT SaliencePush ADS BedRock Roomalizer; do
  echo "== $k $(timeout 300 python tools/prof_run.py --plugin $k --streams 4096 --seconds 1 --reps 2 2>&1 | tail -1 | cut -c1-200)" >> gpurun_out/r7_times.log
done
timeout 900 python bench.py --workload cfg5 --steps 2 --warmup 3 > gpurun_out/bench_cfg5.json 2> gpurun_out/bench_cfg5.err
cat gpurun_out/tests7.log gpurun_out/r7_times.log; cut -c1-200 gpurun_out/bench_cfg5.json
