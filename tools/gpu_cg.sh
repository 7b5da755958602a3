set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_poisson.py -x -q > gpurun_out/pytest_cg.log 2>&1; tail -15 gpurun_out/pytest_cg.log
timeout 600 python tools/kbench.py cg 256 256 256 > gpurun_out/cg256.log 2>&1; tail -5 gpurun_out/cg256.log
timeout 600 python tools/kbench.py cg 512 512 512 > gpurun_out/cg512.log 2>&1; tail -5 gpurun_out/cg512.log
timeout 600 python tools/kbench.py cg 512 512 512 1e-3 > gpurun_out/cg512_1e3.log 2>&1; tail -5 gpurun_out/cg512_1e3.log
ncu --set full --clock-control none --import-source on -k regex:'k_cg_apply|k_cg_update|k_cg_direction|k_cg_finish' -s 40 -c 5 -o gpurun_out/prof_cg2_r01 python tools/kbench.py cg 256 256 256 > gpurun_out/ncu5.log 2>&1
tail -2 gpurun_out/ncu5.log
