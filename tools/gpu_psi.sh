#!/usr/bin/env bash
# psi-only scalar sweeps: fused GPU tests, microbench at 256^3 / 512^3, ncu of one psi-only sweep at 512^3.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_fused.py tests/test_gpu_slab.py -m gpu -x -q > gpurun_out/psi_pytest.log 2>&1
echo "pytest rc=$?"; tail -n 8 gpurun_out/psi_pytest.log
timeout 300 python tools/kbench.py ssweep 256 256 256
timeout 300 python tools/kbench.py ssweep 512 512 512
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_mag_sweep_psi_tiled' -s 6 -c 1 -o gpurun_out/prof_psi_r02 -f \
  python tools/kbench.py ssweep 512 512 512 10 > gpurun_out/psi_ncu.log 2>&1
ncu -i gpurun_out/prof_psi_r02.ncu-rep --page raw --csv > gpurun_out/prof_psi_r02.raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/prof_psi_r02.raw.csv | head -20
