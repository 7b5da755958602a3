# ncu --set full of the extension-tier kernels: one launch per configuration (kbench "ext D H W 1" takes one warm
# and one timed step per configuration; every second launch is captured).  The report stays on the box; the raw
# csv page comes back.
set -x
mkdir -p gpurun_out
timeout 600 python tools/kbench.py ext 256 256 256 1 > gpurun_out/ext256_1step.log 2>&1 && \
ncu --set full --clock-control none -k regex:'k_ext_stream_collide|k_stream_collide' -c 44 -o /tmp/prof_ext_r01 python tools/kbench.py ext 256 256 256 1 > gpurun_out/ncu_ext.log 2>&1
ncu -i /tmp/prof_ext_r01.ncu-rep --page raw --csv > gpurun_out/prof_ext_r01.raw.csv 2>/dev/null
ls -la /tmp/prof_ext_r01.ncu-rep gpurun_out/prof_ext_r01.raw.csv
timeout 600 python -m pytest tests/test_gpu_ext.py -q -k "shan_chen" 2>&1 | tail -5
