mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_easyexpander_gpu.py tests/test_bounce.py tests/test_chain_gpu.py -m gpu -q -s 2>&1 | grep -v "^$" | tail -12
TAG=r02y STEPS=4 BENCH_ARGS="--no-cpu --no-e2e" tools/gpu_run.sh wl:cfg3 > /dev/null
python - <<PY
import json; d=json.loads(open('gpurun_out/r02y_cfg3.json').read().strip().splitlines()[-1]); f=d['f64']; print('f64', f['ms_per_step'], f['roofline']['kernel_ms_per_launch'], f['roofline']['frac'], f['parity_spot_check'])
PY
