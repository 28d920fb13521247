mkdir -p gpurun_out
TAG=r02A PYTEST_ARGS="--maxfail=6" tools/gpu_run.sh tests
for w in erbtilt specstab modtilt; do TAG=r02A STEPS=3 BENCH_ARGS="--no-cpu --no-e2e --no-f64" tools/gpu_run.sh wl:$w > /dev/null; python - <<PY
import json; d=json.loads(open('gpurun_out/r02A_$w.json').read().strip().splitlines()[-1]); r=d['roofline']; print('$w', round(d['ms_per_step'],2), r['kernel_ms_per_launch'], round(r['frac'],4), d['parity_spot_check']['residual_dbfs_rms'])
PY
done
