# One GPU-box call for the N4 extension tier: GPU tests, then the single-phase microbenches.
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; tail -25 gpurun_out/pytest_gpu.log | cut -c1-300
timeout 600 python tools/kbench.py ext 512 512 512 > gpurun_out/ext512.log 2>&1; cat gpurun_out/ext512.log | tail -24
