mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -5
for v in 1 2; do LBMFF_SWEEP_VARIANT=$v python tools/kbench.py sweep 512 512 512; done
LBMFF_SWEEP_VARIANT=1 python tools/kbench.py sweep 256 256 256
LBMFF_SWEEP_VARIANT=1 python tools/kbench.py sweep 130 66 130
