python -m pytest tests/test_gpu_fused.py -m gpu -q -x 2>&1 | tail -3
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-fast-path | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print(d['value'], d['ms_per_step'], d['e2e'], d['clocks'])"
