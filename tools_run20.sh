mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_e2e.json 2> gpurun_out/bench_e2e.err; python - <<PY
import json
d=json.load(open("gpurun_out/bench_e2e.json"))
print(d["value"], d["ms_per_step"], d["roofline"], d["e2e"], d["fast_path"]["value"])
PY
tail -3 gpurun_out/bench_e2e.err
