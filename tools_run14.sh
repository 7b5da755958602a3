set -x
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --no-e2e > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; tail -c 2800 gpurun_out/bench_n2.json; tail -5 gpurun_out/bench_n2.err
TORCH_NCCL_HIGH_PRIORITY=0 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3 --no-e2e --no-fast-path > gpurun_out/bench_n2_lowprio.json 2> gpurun_out/bench_n2_lowprio.err; tail -c 1200 gpurun_out/bench_n2_lowprio.json; tail -5 gpurun_out/bench_n2_lowprio.err
