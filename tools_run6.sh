python -m pytest tests -m gpu -q > gpurun_out/gpu_tests.log 2>&1; tail -3 gpurun_out/gpu_tests.log; grep -E "^E  .*(rel-L2|Error)" gpurun_out/gpu_tests.log | head
python bench.py --steps 5 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/bench_faithful.json 2> gpurun_out/bench_faithful.err; cut -c1-1500 gpurun_out/bench_faithful.json; tail -5 gpurun_out/bench_faithful.err
python bench.py --res 256 256 256 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r01.csv python bench.py --res 256 256 256 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu1.log 2>&1
