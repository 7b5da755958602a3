#!/usr/bin/env bash
# End-of-round check on one GPU: smoke(), the whole gpu suite (-x, as the driver runs it), the default bench line.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/final_smoke.log
python -m pytest tests -m gpu -x -q > gpurun_out/final_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/final_pytest.log
python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench rc=$?"; python tools/show_bench.py gpurun_out/final_bench.json
