mkdir -p gpurun_out
export LBMFF_SWEEP_PAIR=1
python bench.py --res 256 256 256 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-fast-path > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_mag_sweep_pair' -s 16 -c 1 -o gpurun_out/prof_pair_r01 python bench.py --res 256 256 256 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-fast-path > gpurun_out/ncu5.log 2>&1
tail -2 gpurun_out/ncu5.log
