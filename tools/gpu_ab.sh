# same-box A/B of compile-time switches: bench.py with the tree's build, then once per argument after a rebuild of
# fused_hcz.cu with that nvcc flag, then the tree's build again.   bash tools/gpu_ab.sh -DFOO=1 -DFOO=2
set -x
mkdir -p gpurun_out
B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-path"
$B > gpurun_out/ab_base1.json 2>/dev/null; python tools/show_bench.py gpurun_out/ab_base1.json
n=0
for flag in "$@"; do
  n=$((n+1))
  touch lbm_ferrofluid_b200/csrc/fused_hcz.cu; LBMFF_NVCC_EXTRA="$flag" python -m lbm_ferrofluid_b200.build > /dev/null
  echo "variant $n: $flag"
  $B > gpurun_out/ab_v$n.json 2>/dev/null; python tools/show_bench.py gpurun_out/ab_v$n.json
done
touch lbm_ferrofluid_b200/csrc/fused_hcz.cu; python -m lbm_ferrofluid_b200.build > /dev/null
$B > gpurun_out/ab_base2.json 2>/dev/null; python tools/show_bench.py gpurun_out/ab_base2.json
