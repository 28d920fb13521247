python -m pytest tests/test_easyexpander_gpu.py tests/test_hostpath_gpu.py -m gpu -q -x -s 2>&1 | tail -25 > gpurun_out/r02i_tests.log; tail -12 gpurun_out/r02i_tests.log
TAG=r02i STEPS=5 BENCH_ARGS="--no-cpu --no-e2e --no-f64" tools/gpu_run.sh wl:cfg3
TAG=r02i2 STEPS=5 BENCH_ARGS="--no-cpu --no-e2e --no-f64 --option scalar_f32=2" tools/gpu_run.sh wl:cfg3
