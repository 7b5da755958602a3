# Same-box A/B of the halo transport of the 30 magnetic sweeps: NCCL point-to-point vs fused push (symmetric memory).
#   gpurun --gpus N --timeout 1800 -- 'bash tools/gpu_push_ab.sh N'
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29566"
B="bench.py --gpus $N --steps 10 --warmup 3 --no-e2e --no-fast-path --no-cpu-baseline --no-configs --no-ref-eager"
for rep in 1 2; do for push in 0 1; do
  LBMFF_SLAB_PUSH=$push $TR $B > gpurun_out/push_ab_n${N}_p${push}_$rep.json 2> gpurun_out/push_ab_n${N}_p${push}_$rep.err
  echo "push=$push rep=$rep rc=$?"; python tools/show_bench.py gpurun_out/push_ab_n${N}_p${push}_$rep.json | head -1
done; done
