set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_iso_surface.py tests/test_gpu_render.py tests/test_gpu_poisson.py -x -q -s > gpurun_out/pytest_n3.log 2>&1; tail -25 gpurun_out/pytest_n3.log
timeout 600 python tools/kbench.py render 256 > gpurun_out/render256.log 2>&1; tail -3 gpurun_out/render256.log
timeout 600 python tools/kbench.py render 512 > gpurun_out/render512.log 2>&1; tail -3 gpurun_out/render512.log
timeout 600 python tools/kbench.py cg 256 256 256 > gpurun_out/cg256.log 2>&1; tail -4 gpurun_out/cg256.log
timeout 600 python tools/kbench.py cg 512 512 512 > gpurun_out/cg512.log 2>&1; tail -4 gpurun_out/cg512.log
