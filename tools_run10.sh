set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/gpu_tests.log 2>&1; tail -8 gpurun_out/gpu_tests.log
for v in 1 4; do
LBMFF_SWEEP_VARIANT=$v python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_v$v.json 2> gpurun_out/bench_v$v.err; python - <<PY
import json
d=json.load(open("gpurun_out/bench_v$v.json"))
print("variant $v", d["value"], d["ms_per_step"], d["roofline"]["avg_launch_ms"], d["roofline"]["frac"], d["roofline"]["share_of_step"])
PY
tail -3 gpurun_out/bench_v$v.err
done
