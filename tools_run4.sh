set -x
python -m pytest tests -m gpu -q > gpurun_out/gpu_tests.log 2>&1; tail -3 gpurun_out/gpu_tests.log
python bench.py --steps 5 --warmup 2 > gpurun_out/bench_faithful.json 2> gpurun_out/bench_faithful.err; tail -c 2500 gpurun_out/bench_faithful.json; tail -5 gpurun_out/bench_faithful.err
python bench.py --steps 5 --warmup 2 --magnetic-path scalar --no-cpu-baseline > gpurun_out/bench_scalar.json 2> gpurun_out/bench_scalar.err; tail -c 2500 gpurun_out/bench_scalar.json; tail -5 gpurun_out/bench_scalar.err
python bench.py --res 256 256 256 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r01.csv python bench.py --res 256 256 256 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_mag_sweep|k_hcz_collide_stream' -s 35 -c 3 -o gpurun_out/prof_r01 python bench.py --res 256 256 256 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/ncu2.log
