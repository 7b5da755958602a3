set -x
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; tail -c 2500 gpurun_out/bench_n2.json; tail -15 gpurun_out/bench_n2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3 --magnetic-path scalar --no-e2e > gpurun_out/bench_n2_scalar.json 2> gpurun_out/bench_n2_scalar.err; tail -c 1500 gpurun_out/bench_n2_scalar.json; tail -5 gpurun_out/bench_n2_scalar.err
