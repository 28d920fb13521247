mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zfaust_leaves_gpu.py tests/test_chain_gpu.py -m gpu -q -x 2>&1 | tail -2
CMD="python tools/prof_run.py --plugin ClickBeGoneSG --streams 4096 --seconds 2 --reps 2"
$CMD > gpurun_out/r02_cbg_plain.log 2>&1 || exit 1; tail -1 gpurun_out/r02_cbg_plain.log
ncu --set full --clock-control none --import-source on -k regex:cbg_tiled -s 1 -c 1 -o gpurun_out/r02_cbg_tiled -f $CMD > gpurun_out/r02_cbg_ncu.log 2>&1; tail -1 gpurun_out/r02_cbg_ncu.log
