set -x
mkdir -p gpurun_out
show() { python tools/show_bench.py $1; }
B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-path"
for bx in 32 64 128 256; do
LBMFF_HCZ_BX=$bx $B > gpurun_out/pf_bx$bx.json 2>/dev/null; show gpurun_out/pf_bx$bx.json
done
LBMFF_HCZ_BX=32 LBMFF_HCZ_KZ=32 $B > gpurun_out/pf_bx32kz32.json 2>/dev/null; show gpurun_out/pf_bx32kz32.json
