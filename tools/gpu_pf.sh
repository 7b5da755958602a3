set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fused.py tests/test_gpu_slab.py -m gpu -x -q 2>&1 | tail -3
B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-path"
for i in 1 2; do
LBMFF_HCZ_PREFETCH=1 $B > gpurun_out/pf_on$i.json 2>/dev/null; python tools/show_bench.py gpurun_out/pf_on$i.json
LBMFF_HCZ_PREFETCH=0 $B > gpurun_out/pf_off$i.json 2>/dev/null; python tools/show_bench.py gpurun_out/pf_off$i.json
done
