#!/bin/bash
# tools/gpu_final.sh -- one GPU session with the round's final kernels: smoke, all GPU tests, the bench lines of every
# workload (+ the reference arm), filtered launch lists of the benched commands, full ncu captures of the dominant
# kernels.  Outputs land in gpurun_out/ (scratch); tools/make_profiles.py turns them into profiles/.
mkdir -p gpurun_out
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -5 > gpurun_out/tests.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_cfg2.json 2> gpurun_out/bench_ref_cfg2.err
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_cfg2.json 2> gpurun_out/bench_cfg2.err
python bench.py --workload cfg3 --steps 3 --warmup 3 > gpurun_out/bench_cfg3.json 2> gpurun_out/bench_cfg3.err
python bench.py --workload cfg4 --steps 3 --warmup 3 > gpurun_out/bench_cfg4.json 2> gpurun_out/bench_cfg4.err
timeout 900 python bench.py --workload cfg5 --steps 3 --warmup 3 > gpurun_out/bench_cfg5.json 2> gpurun_out/bench_cfg5.err
rm -f gpurun_out/gen_times.log
for k in RTT SOMA ATTACK ERBTilt SpectralStabilizer DDT DPT SaliencePush ADS BedRock Roomalizer; do
  echo "== $k $(timeout 300 python tools/prof_run.py --plugin $k --streams 4096 --seconds 1 --reps 2 2>&1 | tail -1 | cut -c1-200)" >> gpurun_out/gen_times.log
done
K='regex:cascade8|easyexpander|stft_mask|jsfx_|ee_param'
for w in cfg2 cfg3 cfg4; do
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 400 --csv --log-file gpurun_out/launches_$w.csv \
    python bench.py --workload $w --steps 2 --warmup 3 --no-cpu --no-f64 > gpurun_out/ncu_$w.log 2>&1
done
timeout 500 ncu --set full --clock-control none --import-source on -k regex:easyexpander_split -c 1 -f -o gpurun_out/r01_ee_v8 \
  python tools/prof_run.py --plugin EasyExpander --streams 4096 --seconds 2 --reps 1 > gpurun_out/ncu_ee.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:stft_mask -c 1 -f -o gpurun_out/r01_stft_v4 \
  python tools/prof_run.py --plugin stft_mask2048 --streams 512 --seconds 4 --reps 1 > gpurun_out/ncu_stft.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:jsfx_process_kernel -c 1 -f -o gpurun_out/r01_gen_rtt_v6 \
  python tools/prof_run.py --plugin RTT --streams 4096 --seconds 1 --reps 1 > gpurun_out/ncu_rtt.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:cascade8_x2_kernel -c 1 -f -o gpurun_out/r01_c8_v3 \
  python tools/prof_run.py --plugin cascade8 --streams 1024 --seconds 10 --reps 1 > gpurun_out/ncu_c8.log 2>&1
cat gpurun_out/smoke.log gpurun_out/tests.log gpurun_out/gen_times.log; for f in gpurun_out/bench_*.json; do echo "== $f"; cut -c1-300 $f; done; tail -3 gpurun_out/bench_cfg5.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:jsfx_process_kernel -c 1 -f -o gpurun_out/r01_gen_ss_v4 \
  python tools/prof_run.py --plugin SpectralStabilizer --streams 4096 --seconds 0.5 --reps 1 > gpurun_out/ncu_ss.log 2>&1
