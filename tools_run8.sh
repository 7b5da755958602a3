set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/gpu_tests.log 2>&1; tail -15 gpurun_out/gpu_tests.log
