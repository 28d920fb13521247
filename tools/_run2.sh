set -u
export TAG=r02b
tools/gpu_run.sh tests
BENCH_ARGS="--no-cpu --no-e2e --seconds 20" tools/gpu_run.sh wl:modtilt
TAG=r02b_scalar BENCH_ARGS="--no-cpu --no-e2e --seconds 20 --option scalar_f32=1" tools/gpu_run.sh wl:modtilt
BENCH_ARGS="--no-cpu --no-e2e --no-f64 --seconds 2" tools/gpu_run.sh wl:erbtilt wl:specstab
BENCH_ARGS="--no-cpu --no-e2e --no-f64 --seconds 2 --streams 8192" TAG=r02b_8k tools/gpu_run.sh wl:erbtilt
python bench.py --workload modtilt --seconds 4 --steps 1 --warmup 3 --no-cpu --no-e2e --no-parity > gpurun_out/plain_mt.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:modtilt_scan -s 2 -c 1 -o gpurun_out/r02_modtilt_scan python bench.py --workload modtilt --seconds 4 --steps 1 --warmup 3 --no-cpu --no-e2e --no-parity > gpurun_out/ncu_mt.log 2>&1
echo ncu rc=$?
