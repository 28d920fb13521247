timeout 300 python -m pytest tests/test_easyexpander_gpu.py -m gpu -q -x 2>&1 | tail -3
for m in 0 32 16 48; do
  echo -n "role_off=$m: "
  timeout 120 python bench.py --workload cfg3 --seconds 10 --steps 3 --warmup 3 --no-cpu --no-e2e --no-f64 --no-parity --option role_off=$m | grep -o '"kernel_ms_per_launch": [0-9.]*'
done
TAG=r02k STEPS=5 BENCH_ARGS="--no-cpu --no-e2e --no-f64" timeout 300 tools/gpu_run.sh wl:cfg3 | grep -o '"ms_per_step": [0-9.]*\|"frac": [0-9.]*\|residual_dbfs_rms": [-0-9.]*'
