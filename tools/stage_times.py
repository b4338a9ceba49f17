"""Print ms/step and per-stage times from a bench.py JSON line on stdin."""
import json, sys
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(sys.argv[1] if len(sys.argv) > 1 else "", "ms/step", d["ms_per_step"], "Mpix/s", d["value"], "e2e", d["e2e"]["value"],
      " | ".join(f"{k} {v['ms']*1000:.0f}" for k, v in d.get("stages", {}).items()))
