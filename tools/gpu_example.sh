set -x
mkdir -p gpurun_out/frames
timeout 900 python -m pytest tests/test_gpu_iso_surface.py tests/test_gpu_render.py -x -q 2>&1 | tail -3
python examples/rosensweig_3d.py --res 130 130 130 --total_steps 3000 --frame_every 250 --mag_strength 100 --perturb 2 --out gpurun_out/ex_long > gpurun_out/example_long.log 2>&1; tail -2 gpurun_out/example_long.log
python examples/rosensweig_3d.py --res 258 258 258 --total_steps 2000 --frame_every 500 --mag_strength 100 --perturb 2 --out gpurun_out/ex_256 > gpurun_out/example_256.log 2>&1; tail -2 gpurun_out/example_256.log
for f in 001 1001 2001 2751; do cp gpurun_out/ex_long/$f.png gpurun_out/frames/long_$f.png; done
cp gpurun_out/ex_256/final_frame.png gpurun_out/frames/ros256_2000.png
rm -rf gpurun_out/ex_long gpurun_out/ex_256
