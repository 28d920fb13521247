set -u
export TAG=r02e
tools/gpu_run.sh tests
BENCH_ARGS="--no-cpu --no-e2e --seconds 4" tools/gpu_run.sh wl:specstab
BENCH_ARGS="--no-cpu --no-e2e" tools/gpu_run.sh wl:cfg5
python bench.py --workload specstab --seconds 2 --steps 1 --warmup 3 --no-cpu --no-e2e --no-parity --no-f64 > gpurun_out/plain_ss.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:specstab_kernel -s 2 -c 1 -o gpurun_out/r02_specstab python bench.py --workload specstab --seconds 2 --steps 1 --warmup 3 --no-cpu --no-e2e --no-parity --no-f64 > gpurun_out/ncu_ss.log 2>&1
echo ncu rc=$?
