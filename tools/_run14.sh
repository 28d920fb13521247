mkdir -p gpurun_out
for o in "sched=5" "sched=0"; do
python bench.py --workload cfg5 --steps 6 --warmup 3 --no-cpu --no-e2e --option $o > gpurun_out/r02v_cfg5.json 2>gpurun_out/r02v_cfg5.err; python - <<PY
import json; d=json.loads(open('gpurun_out/r02v_cfg5.json').read().strip().splitlines()[-1]); print('$o', round(d['ms_per_step'],1), max(v['residual_dbfs_rms'] for v in d['parity_spot_check'].values()))
PY
done
python bench.py --workload cfg5 --steps 6 --warmup 3 --no-cpu --no-e2e > gpurun_out/r02v_cfg5.json 2>gpurun_out/r02v_cfg5.err; python - <<PY
import json; d=json.loads(open('gpurun_out/r02v_cfg5.json').read().strip().splitlines()[-1]); print('default', round(d['ms_per_step'],1), d['config']['parallelism'])
PY
