mkdir -p gpurun_out
for bx in 256 128 64 32; do
LBMFF_COLLIDE_BX=$bx python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-path > gpurun_out/bench_bx$bx.json 2> gpurun_out/bench_bx$bx.err; python - <<PY
import json
d=json.load(open("gpurun_out/bench_bx$bx.json"))
print("collide bx $bx", d["value"], d["ms_per_step"], d["roofline"]["share_of_step"]["collide_stream"]*d["ms_per_step"])
PY
tail -3 gpurun_out/bench_bx$bx.err
done
