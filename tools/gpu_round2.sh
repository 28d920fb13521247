#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_stft_gpu.py tests/test_easyexpander_gpu.py tests/test_cascade8_gpu.py -m gpu -q 2>&1 | tail -3 > gpurun_out/tests2.log
python bench.py --workload cfg4 --steps 3 --warmup 3 > gpurun_out/bench_cfg4.json 2> gpurun_out/bench_cfg4.err
timeout 900 python bench.py --workload cfg5 --steps 2 --warmup 3 > gpurun_out/bench_cfg5.json 2> gpurun_out/bench_cfg5.err
K='regex:cascade8|easyexpander|stft_mask|jsfx_|ee_param'
for w in cfg2 cfg3 cfg4; do
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 400 --csv --log-file gpurun_out/launches_$w.csv \
    python bench.py --workload $w --steps 2 --warmup 3 --no-cpu --no-f64 > gpurun_out/ncu_$w.log 2>&1
done
timeout 500 ncu --set full --clock-control none --import-source on -k regex:cascade8_x2_kernel -c 1 -f -o gpurun_out/r01_c8_v3 \
  python tools/prof_run.py --plugin cascade8 --streams 1024 --seconds 10 --reps 1 > gpurun_out/ncu_c8.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:stft_mask -c 1 -f -o gpurun_out/r01_stft_v4 \
  python tools/prof_run.py --plugin stft_mask2048 --streams 512 --seconds 4 --reps 1 > gpurun_out/ncu_stft.log 2>&1
cat gpurun_out/tests2.log; cut -c1-300 gpurun_out/bench_cfg4.json; cut -c1-200 gpurun_out/bench_cfg5.json; tail -2 gpurun_out/bench_cfg5.err
