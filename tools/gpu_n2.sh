# 2-GPU session: data-plane check, weak-scaling bench line, configs[4] at 1024^3 on two GPUs (compressed streaming).
#   gpurun --gpus 2 --timeout 2400 -- 'bash tools/gpu_n2.sh'
set -x
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
$TR tools/verify_slab.py > gpurun_out/verify_n$N.json 2> gpurun_out/verify_n$N.err; echo "verify rc=$?"; cat gpurun_out/verify_n$N.json; tail -n 5 gpurun_out/verify_n$N.err
$TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n${N}_r02.json 2> gpurun_out/bench_n${N}_r02.err; echo "bench rc=$?"; python tools/show_bench.py gpurun_out/bench_n${N}_r02.json; tail -n 3 gpurun_out/bench_n${N}_r02.err
D=$((1024 / N))
if [ "$N" = "2" ]; then EXTRA="--magnetic-path scalar --streaming compressed --chunk 32"; else EXTRA=""; fi
$TR bench.py --gpus $N --workload two_droplets --res $D 1024 1024 --steps 5 --warmup 2 --reps 1 --no-fast-path $EXTRA > gpurun_out/bench_two_droplets_1024_n${N}_r02.json 2> gpurun_out/bench_two_droplets_1024_n${N}_r02.err; echo "droplets rc=$?"; python tools/show_bench.py gpurun_out/bench_two_droplets_1024_n${N}_r02.json; tail -n 5 gpurun_out/bench_two_droplets_1024_n${N}_r02.err
