set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_iso_surface.py tests/test_gpu_render.py -x -q -s > gpurun_out/pytest_n3.log 2>&1; tail -25 gpurun_out/pytest_n3.log
timeout 600 python tools/kbench.py render 256 > gpurun_out/render256.log 2>&1; tail -3 gpurun_out/render256.log
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_raster_tile|k_bin|k_group_box|k_face_setup|k_scan' --csv --log-file gpurun_out/launches_render_r01.csv python tools/kbench.py render 256 > gpurun_out/ncu7.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_raster_tile|k_bin' -c 3 -o gpurun_out/prof_raster_r01 python tools/kbench.py render 256 > gpurun_out/ncu6.log 2>&1
tail -2 gpurun_out/ncu6.log
