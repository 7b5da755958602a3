#!/usr/bin/env bash
# ncu --set full of the level-0 multigrid smoother and the CG operator at 512^3 (one launch each).
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python tools/kbench.py pcgprof 512 512 512 > gpurun_out/pcgprof_plain.log 2>&1 || { tail -5 gpurun_out/pcgprof_plain.log; exit 1; }
LBMFF_CG_NO_GRAPH=1 timeout 900 ncu --set full --clock-control none --import-source on \
  -k regex:'k_mg0_sweep_tiled' -s 3 -c 2 -o gpurun_out/prof_pcg_r02 -f \
  python tools/kbench.py pcgprof 512 512 512 > gpurun_out/pcg_ncu_full.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/pcg_ncu_full.log
ncu -i gpurun_out/prof_pcg_r02.ncu-rep --page raw --csv > gpurun_out/prof_pcg_r02.raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/prof_pcg_r02.raw.csv 2>/dev/null | head -80
python - <<'PY'
import csv
rows = list(csv.reader(open("gpurun_out/prof_pcg_r02.raw.csv")))
hdr = rows[0]
for row in rows[2:]:
    d = dict(zip(hdr, row))
    st = []
    for k, v in d.items():
        if "issue_stalled" in k and k.endswith("per_issue_active.ratio"):
            try: st.append((float(v), k))
            except ValueError: pass
    st.sort(reverse=True)
    print(d["Kernel Name"][:60], " | ".join(f"{k.split('issue_stalled_')[1].split('_per_')[0]}={v:.2f}" for v, k in st[:7]))
PY
