#!/usr/bin/env bash
# A/B of the tiled level-0 smoother variants at 512^3: cold multigrid-PCG solve, launch list per variant.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_poisson.py -m gpu -x -q > gpurun_out/pcg_pytest.log 2>&1
echo "pytest rc=$?"; tail -n 6 gpurun_out/pcg_pytest.log
for rows in 2; do for zc in 16 32; do
  echo "== rows=$rows zc=$zc"
  LBMFF_MG_ROWS=$rows LBMFF_MG_ZC=$zc timeout 300 python tools/kbench.py pcgcold 512 512 512 2>&1 | grep "pcg-mg cold"
done; done
timeout 600 python tools/kbench.py pcg 256 256 256 > gpurun_out/pcg256.log 2>&1; grep -v history gpurun_out/pcg256.log
timeout 600 python tools/kbench.py pcg 512 512 512 > gpurun_out/pcg512.log 2>&1; grep -v history gpurun_out/pcg512.log
LBMFF_CG_NO_GRAPH=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_mg|k_cg|k_pcg' -c 1500 --csv \
  --log-file gpurun_out/pcg512_launches.csv python tools/kbench.py pcgprof 512 512 512 > gpurun_out/pcg512_ncu.log 2>&1
python tools/launch_summary.py gpurun_out/pcg512_launches.csv | head -24
