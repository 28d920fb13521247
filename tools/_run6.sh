set -u
export TAG=r02f
tools/gpu_run.sh tests
python -m pytest tests -m gpu -q -k "fft or Alias or chain or heap or specstab or leaves" > gpurun_out/r02f_tests2.log 2>&1; tail -4 gpurun_out/r02f_tests2.log
