#!/usr/bin/env bash
# Converged-solver session: the Poisson GPU tests, CG vs multigrid-PCG at 256^3 / 512^3, and an ncu launch list
# (durations only) of the 512^3 run.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_poisson.py -m gpu -x -q > gpurun_out/pcg_pytest.log 2>&1
echo "pytest rc=$?"; tail -n 6 gpurun_out/pcg_pytest.log
timeout 600 python tools/kbench.py pcg 256 256 256 > gpurun_out/pcg256.log 2>&1; cat gpurun_out/pcg256.log
timeout 600 python tools/kbench.py pcg 512 512 512 > gpurun_out/pcg512.log 2>&1; cat gpurun_out/pcg512.log
LBMFF_CG_NO_GRAPH=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_mg|k_cg|k_pcg' -c 1500 --csv \
  --log-file gpurun_out/pcg512_launches.csv python tools/kbench.py pcgprof 512 512 512 > gpurun_out/pcg512_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/pcg512_ncu.log
python tools/launch_summary.py gpurun_out/pcg512_launches.csv 2>/dev/null | head -40
