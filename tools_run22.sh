mkdir -p gpurun_out
LBMFF_COLLIDE_VARIANT=1 python -m pytest tests/test_gpu_fused.py tests/test_gpu_slab.py -m gpu -q -x 2>&1 | tail -3
for v in 0 1; do
LBMFF_COLLIDE_VARIANT=$v python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-path > gpurun_out/bench_c$v.json 2> gpurun_out/bench_c$v.err; python - <<PY
import json
d=json.load(open("gpurun_out/bench_c$v.json"))
print("collide variant $v", d["value"], d["ms_per_step"], d["roofline"]["share_of_step"]["collide_stream"]*d["ms_per_step"])
PY
tail -3 gpurun_out/bench_c$v.err
done
