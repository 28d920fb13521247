mkdir -p gpurun_out
TAG=r02r PYTEST_ARGS="--maxfail=5" tools/gpu_run.sh tests
# launch list of the benched command (after it exited 0 without ncu in this same call)
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r02r_plain.json 2> gpurun_out/r02r_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/r02_bench_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r02r_ncu.json 2> gpurun_out/r02r_ncu.err
echo "ncu rc=$?"; wc -l gpurun_out/r02_bench_launches.csv
