export LBMFF_SWEEP_PAIR=2
timeout 300 python -m pytest tests/test_gpu_fused.py -m gpu -q -x 2>&1 | tail -6
timeout 120 python tools/kbench.py sweep 512 512 512 | sed "s/^/pair=2 /"
timeout 120 python tools/kbench.py sweep 256 256 256 | sed "s/^/pair=2 /"
