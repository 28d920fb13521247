#!/bin/bash
# four-warp CTAs in the translated leaves: parity, solo timing, scaling with the stream count, cfg5
mkdir -p gpurun_out; rm -f gpurun_out/r8_times.log
timeout 1200 python -m pytest tests/test_generated_gpu.py tests/test_bounce.py -m gpu -q 2>&1 | tail -4 > gpurun_out/tests8.log
for k in RTT SOMA ATTACK ERBTilt SpectralStabilizer DPT SaliencePush ADS Roomalizer; do
  echo "== $k $(timeout 300 python tools/prof_run.py --plugin $k --streams 4096 --seconds 1 --reps 2 2>&1 | tail -1 | cut -c1-200)" >> gpurun_out/r8_times.log
done
for k in SpectralStabilizer RTT; do
  for n in 8192 16384 37888; do
    echo "== $k $n $(timeout 300 python tools/prof_run.py --plugin $k --streams $n --seconds 0.5 --reps 2 2>&1 | tail -1 | cut -c1-200)" >> gpurun_out/r8_times.log
  done
done
timeout 900 python bench.py --workload cfg5 --steps 3 --warmup 3 > gpurun_out/bench_cfg5.json 2> gpurun_out/bench_cfg5.err
cat gpurun_out/tests8.log gpurun_out/r8_times.log; cut -c1-200 gpurun_out/bench_cfg5.json; tail -2 gpurun_out/bench_cfg5.err
