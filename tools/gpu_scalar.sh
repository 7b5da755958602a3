set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fused.py tests/test_gpu_slab.py -x -q 2>&1 | tail -3
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-path --magnetic-path scalar > gpurun_out/bench_scalar.json 2> gpurun_out/bench_scalar.err
python - <<PY
import json
b=json.loads([l for l in open('gpurun_out/bench_scalar.json') if l.startswith('{')][-1])
sh=b['roofline']['share_of_step']
print('scalar path MLUPS', round(b['value'],1), 'ms', round(b['ms_per_step'],2), 'sweeps ms', round(sh['sweeps']*b['ms_per_step'],2), 'per sweep', round(sh['sweeps']*b['ms_per_step']/30,3), 'cs ms', round(sh['collide_stream']*b['ms_per_step'],2))
PY
