mkdir -p gpurun_out
for p in RED VAR; do
CMD="python tools/prof_run.py --plugin $p --streams 5041 --seconds 1 --reps 2"
$CMD 2>&1 | tail -1
lc=$(echo $p | tr A-Z a-z)
ncu --set full --clock-control none --import-source on -k regex:${lc}_tiled -s 1 -c 1 -o gpurun_out/r02_${lc}_tiled -f $CMD > gpurun_out/r02_${lc}_ncu.log 2>&1; tail -1 gpurun_out/r02_${lc}_ncu.log
done
