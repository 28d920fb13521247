#!/bin/bash
# tools/gpu_round13.sh -- ModTilt: frames per straight-line run (8 / 16 / 32) at 4096 and 65 536 streams.
# --warps = streams per warp + 100 x frames per run (0 = 16).
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zfaust_modtilt_gpu.py -m gpu -q 2>&1 | tail -3 > gpurun_out/tests_modtilt13.log
rm -f gpurun_out/modtilt_run_sweep.log
for cfg in "4096 10 408" "4096 10 808" "65536 2 432" "65536 2 832" "16384 4 816" "16384 4 832"; do
  set -- $cfg
  echo "== streams $1 seconds $2 option $3: $(timeout 300 python tools/prof_run.py --plugin ModTilt --streams $1 --seconds $2 --reps 3 --warps $3 2>&1 | tail -1 | cut -c1-300)" >> gpurun_out/modtilt_run_sweep.log
done
cat gpurun_out/tests_modtilt13.log gpurun_out/modtilt_run_sweep.log
