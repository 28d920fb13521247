set -u
export TAG=r02d
tools/gpu_run.sh tests
python -m pytest tests -m gpu -q -k "faust_leaves or erbtilt or quantiser or checkpoint or multi_chunk" > gpurun_out/r02d_tests2.log 2>&1; tail -3 gpurun_out/r02d_tests2.log
BENCH_ARGS="--no-cpu --no-e2e --seconds 4" tools/gpu_run.sh wl:gts wl:var
BENCH_ARGS="--no-cpu --no-e2e --seconds 2" tools/gpu_run.sh wl:cfg5
