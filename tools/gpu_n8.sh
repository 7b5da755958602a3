set -x
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; cut -c1-300 gpurun_out/bench_n8.json; tail -3 gpurun_out/bench_n8.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 5 --warmup 3 --workload two_droplets --res 128 1024 1024 > gpurun_out/bench_drops_n8.json 2> gpurun_out/bench_drops_n8.err; cut -c1-300 gpurun_out/bench_drops_n8.json; tail -3 gpurun_out/bench_drops_n8.err
