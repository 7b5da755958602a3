#!/usr/bin/env bash
# ncu --set full (with source) of k_hcz_moments / k_hcz_fields / k_mag_source at 512^3 (one launch each).
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 1 --reps 1 --no-cpu-baseline --no-e2e --no-fast-path --no-configs --no-ref-eager"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_hcz_moments|k_hcz_fields|k_mag_source' -s 3 -c 3 \
  -o gpurun_out/prof_hcz_512_r02b -f $B > gpurun_out/hcz_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/hcz_ncu.log
ncu -i gpurun_out/prof_hcz_512_r02b.ncu-rep --page raw --csv > gpurun_out/prof_hcz_512_r02b.raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/prof_hcz_512_r02b.raw.csv | grep -v "launch__\|l1tex\|lts__"
