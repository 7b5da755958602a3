#!/usr/bin/env bash
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ops.py tests/test_gpu_fused.py tests/test_gpu_slab.py -m gpu -x -q > gpurun_out/phi_pytest.log 2>&1
echo "pytest rc=$?"; tail -n 5 gpurun_out/phi_pytest.log
python bench.py --no-configs --no-ref-eager --no-cpu-baseline > gpurun_out/phi_bench.json 2> gpurun_out/phi_bench.err; echo "bench rc=$?"; python tools/show_bench.py gpurun_out/phi_bench.json
