"""Prints one steady-state step of a timeline_n*_r*.csv.gz (tools/profile_step.py): per stream, busy time and
the big kernels with start offsets.  usage: python tools/timeline_step.py <csv.gz> [min_us]"""
import csv
import gzip
import re
import sys

rows = list(csv.DictReader(gzip.open(sys.argv[1], "rt")))
min_us = float(sys.argv[2]) if len(sys.argv) > 2 else 40.0
for r in rows:
    r["start_us"], r["dur_us"] = float(r["start_us"]), float(r["dur_us"])
marks = [r["start_us"] for r in rows if "interaction_forward" in r["name"]]
s0, s1 = marks[-2], marks[-1]
sel = [r for r in rows if s0 <= r["start_us"] < s1]
print(f"step = {s1 - s0:.0f} us, {len(sel)} launches")
busy = {}
for r in sel:
    busy[r["stream"]] = busy.get(r["stream"], 0.0) + r["dur_us"]
print("busy us per stream:", {k: round(v) for k, v in busy.items()})
# union of busy intervals over all streams -> idle time of the device
iv = sorted((r["start_us"], r["start_us"] + r["dur_us"]) for r in sel)
tot, cur_s, cur_e = 0.0, None, None
for a, b in iv:
    if cur_e is None or a > cur_e:
        if cur_e is not None:
            tot += cur_e - cur_s
        cur_s, cur_e = a, b
    else:
        cur_e = max(cur_e, b)
tot += (cur_e - cur_s) if cur_e is not None else 0
print(f"device busy (union) {tot:.0f} us, idle {s1 - s0 - tot:.0f} us")
for r in sel:
    if r["dur_us"] >= min_us:
        n = re.sub(r"void |at::native::|b200::|\(anonymous namespace\)::", "", r["name"])[:64]
        print(f"{r['start_us'] - s0:8.0f} {r['dur_us']:7.0f} s{r['stream']:>3} q{float(r.get('queued_us') or 0):7.0f} {n}")
