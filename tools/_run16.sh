mkdir -p gpurun_out
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02z_bench.json 2> gpurun_out/r02z_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r02z_bench.err
python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02z_ref.json 2> gpurun_out/r02z_ref.err; echo "ref rc=$?"
TAG=r02z tools/gpu_run.sh smoke | tail -3
