"""Prints value / ms per step / per-kernel means of a bench.py output file (one JSON line among log lines)."""
import json
import sys

for line in open(sys.argv[1]):
    if line.startswith("{"):
        d = json.loads(line)
        print(d.get("n_gpus"), round(d["value"]), round(d["ms_per_step"], 3),
              {k: round(v["mean_ms_per_call"], 3) for k, v in d.get("kernels", {}).items()})
