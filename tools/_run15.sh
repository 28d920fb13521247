timeout 600 python -m pytest tests/test_easyexpander_gpu.py -m gpu -q -x 2>&1 | tail -2
TAG=r02D STEPS=5 BENCH_ARGS="--no-cpu --no-e2e --no-f64" tools/gpu_run.sh wl:cfg3 > /dev/null
python - <<PY
import json; d=json.loads(open('gpurun_out/r02D_cfg3.json').read().strip().splitlines()[-1]); r=d['roofline']; print('default', d['ms_per_step'], r['kernel_ms_per_launch'], r['frac'], d['parity_spot_check']['residual_dbfs_rms'])
PY
