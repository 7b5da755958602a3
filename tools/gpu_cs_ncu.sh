#!/usr/bin/env bash
# ncu --set full (with source) of one k_hcz_collide_stream_tma launch at 512^3.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
KB_TMA=1 KB_KZ=16 timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_hcz_collide_stream' -s 3 -c 1 \
  -o gpurun_out/prof_cs_512_r02b -f python tools/kbench.py cs 512 512 512 > gpurun_out/cs_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/cs_ncu.log
ncu -i gpurun_out/prof_cs_512_r02b.ncu-rep --page raw --csv > gpurun_out/prof_cs_512_r02b.raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/prof_cs_512_r02b.raw.csv | head -20
