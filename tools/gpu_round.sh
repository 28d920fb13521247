#!/bin/bash
# tools/gpu_round.sh -- one GPU session: tests, the bench lines of every workload, launch lists and full ncu captures.
# Outputs land in gpurun_out/ (scratch); summaries are copied to profiles/ by hand.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -5 > gpurun_out/tests.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_cfg2.json 2> gpurun_out/bench_ref_cfg2.err
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_cfg2.json 2> gpurun_out/bench_cfg2.err
python bench.py --workload cfg3 --steps 3 --warmup 3 > gpurun_out/bench_cfg3.json 2> gpurun_out/bench_cfg3.err
python bench.py --workload cfg4 --steps 3 --warmup 3 > gpurun_out/bench_cfg4.json 2> gpurun_out/bench_cfg4.err
timeout 900 python bench.py --workload cfg5 --steps 2 --warmup 3 > gpurun_out/bench_cfg5.json 2> gpurun_out/bench_cfg5.err
# launch lists (same commands, short): per-launch times are cold-cache and serialised; only the kernel's share is used
for w in cfg2 cfg3 cfg4; do
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$w.csv \
    python bench.py --workload $w --steps 2 --warmup 3 --no-cpu --no-f64 > gpurun_out/ncu_$w.log 2>&1
done
# full captures of the dominant kernels (one launch each, short renders)
timeout 500 ncu --set full --clock-control none --import-source on -k regex:easyexpander_kernel -c 1 -f -o gpurun_out/r01_ee_v5 \
  python tools/prof_run.py --plugin EasyExpander --streams 4096 --seconds 2 --reps 1 > gpurun_out/ncu_ee.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:stft_mask -c 1 -f -o gpurun_out/r01_stft_v3 \
  python tools/prof_run.py --plugin stft_mask2048 --streams 512 --seconds 4 --reps 1 > gpurun_out/ncu_stft.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:jsfx_process_kernel -c 1 -f -o gpurun_out/r01_gen_rtt \
  python tools/prof_run.py --plugin RTT --streams 4096 --seconds 1 --reps 1 > gpurun_out/ncu_rtt.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:cascade8_x2 -c 1 -f -o gpurun_out/r01_c8_v3 \
  python tools/prof_run.py --plugin cascade8 --streams 1024 --seconds 10 --reps 1 > gpurun_out/ncu_c8.log 2>&1
cat gpurun_out/tests.log; for f in gpurun_out/bench_*.json; do echo "== $f"; cut -c1-400 $f; done; tail -3 gpurun_out/bench_cfg5.err
