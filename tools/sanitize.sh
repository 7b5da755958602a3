#!/usr/bin/env bash
# compute-sanitizer over small grids (SURVEY.md section 5 / VERDICT r1 item 9): memcheck and racecheck of the smoke
# case (fused 3-D Rosensweig step, both magnetic paths) and of one slab virtual-rank case.  Run on the GPU box:
#     gpurun --timeout 1500 -- 'bash tools/sanitize.sh'
# Logs land in gpurun_out/sanitize_<tool>_<case>.log; the summaries are copied to profiles/ by hand.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
CS=${COMPUTE_SANITIZER:-/usr/local/cuda/bin/compute-sanitizer}
rc=0
for tool in memcheck racecheck; do
  $CS --tool $tool --print-limit 20 --error-exitcode 9 \
      python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/sanitize_${tool}_smoke.log 2>&1 || rc=1
  tail -n 4 gpurun_out/sanitize_${tool}_smoke.log
  $CS --tool $tool --print-limit 20 --error-exitcode 9 \
      python -m pytest tests/test_gpu_slab.py -q -x -k "virtual_ranks_bit_identical and P2 and walls_open0" \
      > gpurun_out/sanitize_${tool}_slab.log 2>&1 || rc=1
  tail -n 4 gpurun_out/sanitize_${tool}_slab.log
done
exit $rc
