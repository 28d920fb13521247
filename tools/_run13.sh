mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_easyexpander_gpu.py tests/test_generated_gpu.py tests/test_zfaust_leaves_gpu.py -m gpu -q -x -k "detector_filter or streams_per_warp" --tb=short 2>&1 | tail -12
for o in sched=0 sched=3 sched=0 sched=3; do
python bench.py --workload cfg5 --steps 6 --warmup 3 --no-cpu --no-e2e --no-parity --option $o > gpurun_out/r02s_cfg5.json 2>gpurun_out/r02s_cfg5.err; python - <<PY
import json; d=json.loads(open('gpurun_out/r02s_cfg5.json').read().strip().splitlines()[-1]); print('$o', round(d['ms_per_step'],1), {k:round(v,1) for k,v in d['roofline']['per_leaf_kernel_ms'].items()})
PY
done
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-f64 --no-parity --subs ,"
$CMD > gpurun_out/r02s_head.json 2> gpurun_out/r02s_head.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_headline_launches.csv $CMD > gpurun_out/r02s_head_ncu.json 2> gpurun_out/r02s_head_ncu.err
echo "ncu rc=$?"; wc -l gpurun_out/r02_headline_launches.csv
