set -u
export TAG=r02c
tools/gpu_run.sh tests
BENCH_ARGS="--no-cpu --no-e2e --seconds 20" tools/gpu_run.sh wl:modtilt
BENCH_ARGS="--no-cpu --no-e2e --seconds 20 --option tile_buffers=2" TAG=r02c_exact tools/gpu_run.sh wl:modtilt
BENCH_ARGS="--no-cpu --no-e2e --seconds 4" tools/gpu_run.sh wl:erbtilt
python bench.py --workload erbtilt --seconds 2 --steps 1 --warmup 3 --no-cpu --no-e2e --no-parity --no-f64 > gpurun_out/plain_erb.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:erbtilt_kernel -s 2 -c 1 -o gpurun_out/r02_erbtilt python bench.py --workload erbtilt --seconds 2 --steps 1 --warmup 3 --no-cpu --no-e2e --no-parity --no-f64 > gpurun_out/ncu_erb.log 2>&1
echo ncu rc=$?
