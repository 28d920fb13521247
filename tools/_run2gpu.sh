mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m pytest tests/test_hostpath_gpu.py tests/test_bounce.py -m gpu -q -k "two_devices or several_gpus or device_list" -rs 2>&1 | tail -6
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02_bench_2gpu.json 2> gpurun_out/r02_bench_2gpu.err; echo "bench2 rc=$?"; tail -3 gpurun_out/r02_bench_2gpu.err; head -c 600 gpurun_out/r02_bench_2gpu.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r02_ref_2gpu.json 2> gpurun_out/r02_ref_2gpu.err; echo "ref2 rc=$?"; head -c 300 gpurun_out/r02_ref_2gpu.json
