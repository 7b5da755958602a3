#!/usr/bin/env bash
# One GPU-box session: the new no-alloc test alone, the whole gpu suite, the converged-solver microbench (CG vs PCG),
# the default bench line.  Logs into gpurun_out/.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r2b_noalloc_alone.log 2>&1
echo "noalloc alone rc=$?"; tail -n 5 gpurun_out/r2b_noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -x -q --durations=15 > gpurun_out/r2b_pytest.log 2>&1
echo "pytest rc=$?"; tail -n 30 gpurun_out/r2b_pytest.log
timeout 600 python tools/kbench.py pcg 256 256 256 > gpurun_out/r2b_cg256.log 2>&1; tail -8 gpurun_out/r2b_cg256.log
timeout 600 python tools/kbench.py pcg 512 512 512 > gpurun_out/r2b_cg512.log 2>&1; tail -8 gpurun_out/r2b_cg512.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err
echo "bench rc=$?"; tail -c 600 gpurun_out/r2b_bench.err; head -c 4000 gpurun_out/r2b_bench.json
