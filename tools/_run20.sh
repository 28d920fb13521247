mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_generated_gpu.py -m gpu -q -x -k "stage_pipeline or SaliencePush or ADS" 2>&1 | tail -6
for p in SaliencePush ADS; do for u in 1 0; do python tools/prof_run.py --plugin $p --streams 5041 --seconds 1 --reps 2 --opt unstaged=$u 2>&1 | tail -1; done; done
