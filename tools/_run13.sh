mkdir -p gpurun_out
for w in 1 2 4; do for nb in 1 2; do
  echo -n "cascade8 warps=$w nbuf=$nb: "
  timeout 120 python bench.py --workload cfg2 --seconds 30 --steps 3 --warmup 3 --no-cpu --no-e2e --no-f64 --no-parity --warps $w --option tile_buffers=$nb 2>&1 | grep -o '"kernel_ms_per_launch": [0-9.]*\|unknown\|error[^"]*' | head -2 | tr '\n' ' '; echo
done; done
for o in sched=0 sched=2 sched=0 sched=2; do
python bench.py --workload cfg5 --steps 6 --warmup 3 --no-cpu --no-e2e --no-parity --option $o > gpurun_out/r02p_cfg5.json 2>gpurun_out/r02p_cfg5.err; python - <<PY
import json; d=json.loads(open('gpurun_out/r02p_cfg5.json').read().strip().splitlines()[-1]); print('$o', round(d['ms_per_step'],1), {k:round(v,1) for k,v in d['roofline']['per_leaf_kernel_ms'].items()})
PY
done
TAG=r02p tools/gpu_run.sh smoke
