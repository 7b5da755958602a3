# same-box A/B of a compile-time switch: bench with the tree's build, rebuild with $1, bench again, rebuild plain
set -x
mkdir -p gpurun_out
show() { python tools/show_bench.py $1; }
B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-path"
$B > gpurun_out/ab_a1.json 2>/dev/null; show gpurun_out/ab_a1.json
touch lbm_ferrofluid_b200/csrc/hcz_fast.cuh; LBMFF_NVCC_EXTRA="$1" python -m lbm_ferrofluid_b200.build > /dev/null
$B > gpurun_out/ab_b1.json 2>/dev/null; show gpurun_out/ab_b1.json
$B > gpurun_out/ab_b2.json 2>/dev/null; show gpurun_out/ab_b2.json
touch lbm_ferrofluid_b200/csrc/hcz_fast.cuh; python -m lbm_ferrofluid_b200.build > /dev/null
$B > gpurun_out/ab_a2.json 2>/dev/null; show gpurun_out/ab_a2.json
