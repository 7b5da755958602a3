# Round-2 evidence capture on the box (512^3, the bench workload):  gpurun --timeout 3000 -- 'bash tools/gpu_profile_r02.sh'
set -x
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 1 --reps 1 --no-cpu-baseline --no-e2e --no-fast-path --no-configs --no-ref-eager"
$B > gpurun_out/prof_plain_r02.json 2> gpurun_out/prof_plain_r02.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_512_r02.csv $B > gpurun_out/ncu1.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_mag_sweep' -s 35 -c 1 -o gpurun_out/prof_sweep_512_r02 $B > gpurun_out/ncu2.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_hcz_collide_stream' -s 1 -c 1 -o gpurun_out/prof_cs_512_r02 $B > gpurun_out/ncu3.log 2>&1
timeout 1200 ncu --set full --clock-control none -k regex:'k_hcz_moments|k_hcz_fields|k_mag_source' -s 3 -c 3 -o gpurun_out/prof_hcz_512_r02 $B > gpurun_out/ncu4.log 2>&1
timeout 1200 ncu --set full --clock-control none -k regex:'k_mag_sweep_scalar' -s 35 -c 1 -o gpurun_out/prof_scalar_512_r02 $B --magnetic-path scalar > gpurun_out/ncu5.log 2>&1
for f in sweep cs hcz scalar; do
  ncu -i gpurun_out/prof_${f}_512_r02.ncu-rep --page raw --csv > gpurun_out/prof_${f}_512_r02.raw.csv 2>/dev/null
done
ncu -i gpurun_out/prof_cs_512_r02.ncu-rep --page source --csv > gpurun_out/prof_cs_512_r02.source.csv 2>/dev/null
ls -la gpurun_out/*.ncu-rep; tail -n 3 gpurun_out/ncu2.log gpurun_out/ncu3.log
