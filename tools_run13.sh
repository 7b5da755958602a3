mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -5
for path in faithful scalar; do
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --magnetic-path $path > gpurun_out/bench_$path.json 2> gpurun_out/bench_$path.err; python - <<PY
import json
d=json.load(open("gpurun_out/bench_$path.json"))
print("$path", d["value"], d["ms_per_step"], d["roofline"]["avg_launch_ms"], d["roofline"]["frac"], d["roofline"]["share_of_step"])
PY
tail -3 gpurun_out/bench_$path.err
done
