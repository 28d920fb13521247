TAG=r02h PYTEST_ARGS="--maxfail=8" tools/gpu_run.sh tests
