for mb in 0 5 6 8; do LBMFF_SWEEP_MB=$mb python tools/kbench.py sweep 512 512 512 | sed "s/^/mb=$mb /"; done
