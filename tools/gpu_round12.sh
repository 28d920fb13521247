#!/bin/bash
# tools/gpu_round12.sh -- GPU session for the Faust leaf ModTilt: parity tests, lane-count / stream-count sweep, bench
# line (+ reference arm), launch list of the benched command, one full ncu capture of the kernel.
mkdir -p gpurun_out
timeout 600 python __graft_entry__.py smoke 2>&1 | tail -6 > gpurun_out/smoke12.log
timeout 600 python -m pytest tests/test_zfaust_modtilt_gpu.py -m gpu -q 2>&1 | tail -15 > gpurun_out/tests_modtilt.log
timeout 900 python bench.py --workload modtilt --steps 3 --warmup 3 > gpurun_out/bench_modtilt.json 2> gpurun_out/bench_modtilt.err
timeout 600 python bench.py --workload modtilt --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_modtilt.json 2> gpurun_out/bench_ref_modtilt.err
rm -f gpurun_out/modtilt_sweep.log
for cfg in "4096 10 8" "4096 10 16" "4096 10 32" "16384 4 0" "16384 4 32" "65536 2 0"; do
  set -- $cfg
  echo "== streams $1 seconds $2 lanes $3: $(timeout 300 python tools/prof_run.py --plugin ModTilt --streams $1 --seconds $2 --reps 3 --warps $3 2>&1 | tail -1 | cut -c1-300)" >> gpurun_out/modtilt_sweep.log
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:modtilt -c 400 --csv --log-file gpurun_out/launches_modtilt.csv \
  python bench.py --workload modtilt --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_modtilt_list.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:modtilt_kernel -c 1 -f -o gpurun_out/r01_modtilt_v2 \
  python tools/prof_run.py --plugin ModTilt --streams 4096 --seconds 2 --reps 1 > gpurun_out/ncu_modtilt.log 2>&1
cat gpurun_out/smoke12.log gpurun_out/tests_modtilt.log gpurun_out/modtilt_sweep.log; cut -c1-1500 gpurun_out/bench_modtilt.json; tail -3 gpurun_out/bench_modtilt.err; cut -c1-400 gpurun_out/bench_ref_modtilt.json
