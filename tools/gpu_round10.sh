#!/bin/bash
# 16-frame tiles for heap leaves: parity, solo timing, cfg5
mkdir -p gpurun_out; rm -f gpurun_out/r10_times.log
timeout 1200 python -m pytest tests/test_generated_gpu.py -m gpu -q 2>&1 | tail -4 > gpurun_out/tests10.log
for k in SOMA ERBTilt SpectralStabilizer DPT Roomalizer; do
  echo "== $k $(timeout 300 python tools/prof_run.py --plugin $k --streams 4096 --seconds 1 --reps 2 2>&1 | tail -1 | cut -c1-200)" >> gpurun_out/r10_times.log
done
timeout 900 python bench.py --workload cfg5 --steps 3 --warmup 3 > gpurun_out/bench_cfg5.json 2> gpurun_out/bench_cfg5.err
cat gpurun_out/tests10.log gpurun_out/r10_times.log; cut -c1-200 gpurun_out/bench_cfg5.json; tail -2 gpurun_out/bench_cfg5.err
