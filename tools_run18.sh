mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -5
for v in 0 1; do
LBMFF_SCALAR_VARIANT=$v python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-path --magnetic-path scalar > gpurun_out/bench_sc$v.json 2> gpurun_out/bench_sc$v.err; python - <<PY
import json
d=json.load(open("gpurun_out/bench_sc$v.json"))
print("scalar variant $v", d["value"], d["ms_per_step"], d["roofline"]["avg_launch_ms"], d["roofline"]["frac"], d["roofline"]["share_of_step"])
PY
tail -3 gpurun_out/bench_sc$v.err
done
