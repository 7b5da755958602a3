#!/usr/bin/env bash
# Tiled CG operator: Poisson GPU tests, CG / PCG at 512^3 with the tiled and the generic operator kernel.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_poisson.py -m gpu -x -q > gpurun_out/cgt_pytest.log 2>&1
echo "pytest rc=$?"; tail -n 5 gpurun_out/cgt_pytest.log
for t in 1 0; do
  echo "== LBMFF_CG_TILED=$t"
  LBMFF_CG_TILED=$t timeout 600 python tools/kbench.py pcg 512 512 512 2>&1 | grep -v history | tee gpurun_out/cgt_pcg512_t$t.log
done
