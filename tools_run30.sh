LBMFF_SWEEP_PAIR=1 python -m pytest tests -m gpu -q -x 2>&1 | tail -6
for v in 0 1; do LBMFF_SWEEP_PAIR=$v python tools/kbench.py sweep 512 512 512 | sed "s/^/pair=$v /"; done
LBMFF_SWEEP_PAIR=1 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-path | python -c "
import json,sys
d=json.loads(sys.stdin.read())
ms=d['ms_per_step']; sh=d['roofline']['share_of_step']
print(d['value'], ms, 'sweeps', sh['sweeps']*ms, 'collide', sh['collide_stream']*ms, 'other', ms*(1-sh['sweeps']-sh['collide_stream']))"
