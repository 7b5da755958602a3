mkdir -p gpurun_out
for v in 1 2 3; do LBMFF_SWEEP_VARIANT=$v python tools/kbench.py sweep 512 512 512; done
for d in 510 511 514; do LBMFF_SWEEP_VARIANT=1 python tools/kbench.py sweep $d 512 512; done
LBMFF_SWEEP_VARIANT=1 python tools/kbench.py sweep 512 512 520
LBMFF_SWEEP_VARIANT=1 python tools/kbench.py sweep 512 520 512
LBMFF_SWEEP_VARIANT=4 python tools/kbench.py sweep 510 512 512
LBMFF_SWEEP_VARIANT=1 python tools/kbench.py sweep 256 256 256
