mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_easyexpander_gpu.py tests/test_hostpath_gpu.py tests/test_chain_gpu.py -m gpu -q -x 2>&1 | tail -3
TAG=r02x STEPS=10 BENCH_ARGS="--no-cpu --no-e2e --no-f64" tools/gpu_run.sh wl:cfg3 > /dev/null
python - <<PY
import json; d=json.loads(open('gpurun_out/r02x_cfg3.json').read().strip().splitlines()[-1]); r=d['roofline']; print(d['ms_per_step'], r['kernel_ms_per_launch'], r['launches_per_step'], r['frac'], d['gpu_launches'], d['parity_spot_check']['residual_dbfs_rms'])
PY
