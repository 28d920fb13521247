#!/bin/bash
# two-GPU check of the torchrun path (streams sharded, no collective) + the tests added after the last full session
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_bounce.py tests/test_generated_gpu.py tests/test_jsfxc.py -q -k "bounce or cfg1" 2>&1 | tail -4 > gpurun_out/tests_new.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench2_cfg2.json 2> gpurun_out/bench2_cfg2.err
timeout 600 $TR bench.py --gpus 2 --impl reference --steps 1 --warmup 1 > gpurun_out/bench2_ref.json 2> gpurun_out/bench2_ref.err
timeout 900 $TR bench.py --gpus 2 --workload cfg5 --steps 2 --warmup 3 > gpurun_out/bench2_cfg5.json 2> gpurun_out/bench2_cfg5.err
timeout 600 $TR bench.py --gpus 2 --workload cfg3 --steps 2 --warmup 3 --no-f64 > gpurun_out/bench2_cfg3.json 2> gpurun_out/bench2_cfg3.err
cat gpurun_out/tests_new.log; for f in gpurun_out/bench2_*.json; do echo "== $f"; cut -c1-260 $f; done; tail -2 gpurun_out/bench2_*.err
