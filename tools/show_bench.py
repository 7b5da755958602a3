"""Print the step / collide-stream / sweep times of a bench.py JSON line (tuning A/B runs)."""
import json
import sys

for path in sys.argv[1:]:
    d = json.load(open(path))
    r = d["roofline"]
    print(path, round(d["value"], 1), "MLUPS", round(d["ms_per_step"], 2), "ms  collide_ms",
          round(r["share_of_step"]["collide_stream"] * d["ms_per_step"], 2), "sweep_ms", round(r["avg_launch_ms"], 3),
          "sm_mhz", d["clocks"]["sm_mhz"])
    for k in ("slab_bit_identical", "exchange"):
        if k in d:
            print("   ", k, d[k])
    if d.get("e2e"):
        print("    e2e", round(d["e2e"]["value"], 1), "fast", round(d["fast_path"]["value"], 1) if d.get("fast_path") else None,
              "state_gb", d["config"].get("state_gb_per_gpu"), "path", d["config"].get("magnetic_path"),
              "streaming", d["config"].get("streaming"))
