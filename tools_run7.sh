set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/gpu_tests.log 2>&1; tail -3 gpurun_out/gpu_tests.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_faithful.json 2> gpurun_out/bench_faithful.err; cut -c1-2500 gpurun_out/bench_faithful.json; tail -5 gpurun_out/bench_faithful.err
python bench.py --res 256 256 256 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_mag_sweep' -s 35 -c 2 -o gpurun_out/prof_sweep_r01 python bench.py --res 256 256 256 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_hcz_collide_stream|k_hcz_moments|k_hcz_fields|k_mag_source' -s 4 -c 4 -o gpurun_out/prof_hcz_r01 python bench.py --res 256 256 256 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu3.log 2>&1
tail -2 gpurun_out/ncu3.log
nproc; free -g | head -2
