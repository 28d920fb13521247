#!/bin/bash
# packed-fp32 EasyExpander band warps + float addressing in translated leaves: parity, timing, captures
mkdir -p gpurun_out; rm -f gpurun_out/r5_times.log
timeout 900 python -m pytest tests/test_easyexpander_gpu.py tests/test_generated_gpu.py -m gpu -q 2>&1 | tail -6 > gpurun_out/tests5.log
echo "== EE $(timeout 300 python tools/prof_run.py --plugin EasyExpander --streams 4096 --seconds 2 --reps 3 2>&1 | tail -1 | cut -c1-200)" >> gpurun_out/r5_times.log
for k in RTT SOMA ERBTilt SpectralStabilizer DDT DPT Roomalizer; do
  echo "== $k $(timeout 300 python tools/prof_run.py --plugin $k --streams 4096 --seconds 1 --reps 2 2>&1 | tail -1 | cut -c1-200)" >> gpurun_out/r5_times.log
done
python bench.py --workload cfg3 --steps 3 --warmup 3 > gpurun_out/bench_cfg3.json 2> gpurun_out/bench_cfg3.err
timeout 500 ncu --set full --clock-control none --import-source on -k regex:easyexpander_split -c 1 -f -o gpurun_out/r01_ee_v7 \
  python tools/prof_run.py --plugin EasyExpander --streams 4096 --seconds 2 --reps 1 > gpurun_out/ncu_ee.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:jsfx_process_kernel -c 1 -f -o gpurun_out/r01_gen_rtt_v2 \
  python tools/prof_run.py --plugin RTT --streams 4096 --seconds 1 --reps 1 > gpurun_out/ncu_rtt.log 2>&1
cat gpurun_out/tests5.log gpurun_out/r5_times.log; cut -c1-400 gpurun_out/bench_cfg3.json; tail -3 gpurun_out/bench_cfg3.err
