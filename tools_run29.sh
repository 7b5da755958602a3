mkdir -p gpurun_out
export NCCL_DEBUG=WARN
N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 2> gpurun_out/n$N.err > gpurun_out/bench_n$N.json; python -c "
import sys,json
for l in open('gpurun_out/bench_n$N.json'):
    if l.startswith('{'):
        d=json.loads(l); print(d['n_gpus'], d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e'] and d['e2e']['value'], d['fast_path'] and d['fast_path']['value'], d['clocks'])
"; grep -m3 "illegal\|Error" gpurun_out/n$N.err; true
