#!/bin/bash
# L1-preferring carve-out for heap leaves + two-phase cfg5: parity, scaling of SpectralStabilizer, cfg5
mkdir -p gpurun_out; rm -f gpurun_out/r11_times.log
timeout 1200 python -m pytest tests/test_generated_gpu.py -m gpu -q 2>&1 | tail -3 > gpurun_out/tests11.log
for n in 4096 8192 16384 37888; do
  echo "== SpectralStabilizer $n $(timeout 300 python tools/prof_run.py --plugin SpectralStabilizer --streams $n --seconds 0.5 --reps 2 2>&1 | tail -1 | cut -c1-200)" >> gpurun_out/r11_times.log
done
for k in ERBTilt SOMA Roomalizer DPT; do
  echo "== $k 4096 $(timeout 300 python tools/prof_run.py --plugin $k --streams 4096 --seconds 1 --reps 2 2>&1 | tail -1 | cut -c1-200)" >> gpurun_out/r11_times.log
done
timeout 900 python bench.py --workload cfg5 --steps 3 --warmup 3 > gpurun_out/bench_cfg5.json 2> gpurun_out/bench_cfg5.err
cat gpurun_out/tests11.log gpurun_out/r11_times.log; cut -c1-200 gpurun_out/bench_cfg5.json; tail -2 gpurun_out/bench_cfg5.err
