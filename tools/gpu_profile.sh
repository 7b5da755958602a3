set -x
mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; cut -c1-3000 gpurun_out/bench_full.json; tail -3 gpurun_out/bench_full.err
python bench.py --res 256 256 256 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-fast-path > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r01.csv python bench.py --res 256 256 256 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-fast-path > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_mag_sweep' -s 35 -c 2 -o gpurun_out/prof_sweep_r01 python bench.py --res 256 256 256 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-fast-path > gpurun_out/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_hcz_collide_stream|k_hcz_moments|k_hcz_fields|k_mag_source' -s 4 -c 4 -o gpurun_out/prof_hcz_r01 python bench.py --res 256 256 256 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-fast-path > gpurun_out/ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_mag_sweep_scalar' -s 35 -c 1 -o gpurun_out/prof_scalar_r01 python bench.py --res 256 256 256 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-fast-path --magnetic-path scalar > gpurun_out/ncu4.log 2>&1
tail -2 gpurun_out/ncu4.log
