#!/bin/bash
# tools/gpu_run.sh STAGE... -- the commands of one gpurun call (run from the repo root ON THE GPU BOX).
# Outputs land in gpurun_out/<tag>_*; nothing here is a bench value when it ran under ncu.
#   tests        python -m pytest tests -m gpu -x -q   (PYTEST_ARGS replaces -x)
#   bench        the default bench line (cfg3 headline + sub-records)
#   ref          the reference arm of the default line
#   latency      bench.py --latency
#   wl:<name>    bench.py --workload <name> (one workload as headline, no sub-records)
set -u
TAG=${TAG:-r02}
mkdir -p gpurun_out
for st in "$@"; do
  case "$st" in
    tests)   python -m pytest tests -m gpu ${PYTEST_ARGS:--x} -q > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/${TAG}_tests.log ;;
    bench)   python bench.py --steps ${STEPS:-10} --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/${TAG}_bench.err; head -c 1500 gpurun_out/${TAG}_bench.json ;;
    ref)     python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_ref.json 2> gpurun_out/${TAG}_ref.err; echo "ref rc=$?"; head -c 600 gpurun_out/${TAG}_ref.json ;;
    latency) python bench.py --latency > gpurun_out/${TAG}_latency.json 2> gpurun_out/${TAG}_latency.err; echo "latency rc=$?"; head -c 800 gpurun_out/${TAG}_latency.json ;;
    wl:*)    w=${st#wl:}; python bench.py --workload $w --steps ${STEPS:-5} --warmup 3 ${BENCH_ARGS:-} > gpurun_out/${TAG}_$w.json 2> gpurun_out/${TAG}_$w.err; echo "$w rc=$?"; tail -2 gpurun_out/${TAG}_$w.err; head -c 1200 gpurun_out/${TAG}_$w.json; echo ;;
    smoke)   python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -8 gpurun_out/${TAG}_smoke.log ;;
    *) echo "unknown stage $st" ;;
  esac
done
