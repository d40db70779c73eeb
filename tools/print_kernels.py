"""reads bench.py's JSON line on stdin and prints step and per-kernel times (tuning aid)"""
import json
import sys

tag = sys.argv[1] if len(sys.argv) > 1 else ""
for line in sys.stdin:
    if not line.startswith("{"):
        continue
    d = json.loads(line)
    print(tag, "step_ms=%.4f" % d["ms_per_step"], "e2e=%.0f" % d["e2e"]["value"],
          {k: round(v["ms_per_step"], 4) for k, v in d["kernels"].items()})
    if "sad" in d:
        s = d["sad"]
        print(tag, "sad step_ms=%.4f" % s["ms_per_step"], "e2e=%.0f" % s["e2e"]["value"],
              {k: round(v["ms_per_step"], 4) for k, v in s["kernels"].items()})
