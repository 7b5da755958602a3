#!/usr/bin/env bash
# One GPU-box session: the gpu test suite, the default bench line and the reference arm, logs into gpurun_out/.
#     gpurun --timeout 2400 -- 'bash tools/gpu_check.sh [tag]'
set -u
cd "$(dirname "$0")/.."
tag=${1:-check}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/${tag}_smi.txt 2>&1
python -m pytest tests -m gpu -q --durations=15 > gpurun_out/${tag}_pytest.log 2>&1
echo "pytest rc=$?"; tail -n 30 gpurun_out/${tag}_pytest.log
python bench.py --steps 10 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
echo "bench rc=$?"; tail -c 600 gpurun_out/${tag}_bench.err; head -c 3000 gpurun_out/${tag}_bench.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2> gpurun_out/${tag}_bench_ref.err
echo "ref rc=$?"; cat gpurun_out/${tag}_bench_ref.json
