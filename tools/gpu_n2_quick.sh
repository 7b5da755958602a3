# 2-GPU sanity session: NCCL slab tests, data-plane check, weak-scaling bench line.
#   gpurun --gpus 2 --timeout 1800 -- 'bash tools/gpu_n2_quick.sh'
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 900 python -m pytest tests/test_gpu_slab.py -m gpu -x -q > gpurun_out/n2_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/n2_pytest.log
$TR tools/verify_slab.py > gpurun_out/verify_n2.json 2> gpurun_out/verify_n2.err; echo "verify rc=$?"; cat gpurun_out/verify_n2.json; tail -n 3 gpurun_out/verify_n2.err
$TR bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2_r02.json 2> gpurun_out/bench_n2_r02.err; echo "bench rc=$?"; python tools/show_bench.py gpurun_out/bench_n2_r02.json; tail -n 3 gpurun_out/bench_n2_r02.err
