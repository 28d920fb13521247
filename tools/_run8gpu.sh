mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l); echo "GPUs: $N"; free -g | head -2; nproc
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r02_bench_${N}gpu.json 2> gpurun_out/r02_bench_${N}gpu.err; echo "bench$N rc=$?"; tail -3 gpurun_out/r02_bench_${N}gpu.err | cut -c1-300; head -c 400 gpurun_out/r02_bench_${N}gpu.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/r02_ref_${N}gpu.json 2> gpurun_out/r02_ref_${N}gpu.err; echo "ref$N rc=$?"; head -c 200 gpurun_out/r02_ref_${N}gpu.json
