# 4-GPU sanity session: data-plane check and the default weak-scaling bench line.
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29577"
$TR tools/verify_slab.py > gpurun_out/verify_n4.json 2> gpurun_out/verify_n4.err; echo "verify rc=$?"; cat gpurun_out/verify_n4.json
$TR bench.py --gpus 4 --steps 10 --warmup 3 > gpurun_out/bench_n4_r02.json 2> gpurun_out/bench_n4_r02.err; echo "bench rc=$?"; python tools/show_bench.py gpurun_out/bench_n4_r02.json; tail -n 3 gpurun_out/bench_n4_r02.err
