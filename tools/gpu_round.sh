# One GPU-box call: GPU test-suite, the converged-solver microbench, then the profile capture of tools/gpu_profile.sh.
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -15 gpurun_out/pytest_gpu.log
timeout 600 python tools/kbench.py cg 256 256 256 > gpurun_out/cg256.log 2>&1; tail -5 gpurun_out/cg256.log
timeout 600 python tools/kbench.py cg 512 512 512 > gpurun_out/cg512.log 2>&1; tail -5 gpurun_out/cg512.log
bash tools/gpu_profile.sh
ncu --set full --clock-control none --import-source on -k regex:'k_cg_apply|k_cg_update|k_cg_direction' -s 30 -c 3 -o gpurun_out/prof_cg_r01 python tools/kbench.py cg 256 256 256 > gpurun_out/ncu5.log 2>&1
tail -2 gpurun_out/ncu5.log
