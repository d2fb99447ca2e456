"""Turns an `ncu --metrics gpu__time_duration.sum --csv` launch list of bench.py into a per-step
markdown table (one training step = the launches between two tbe_forward_kernel launches)."""
import collections
import csv
import re
import sys


def main(path: str, step_index: int = 3) -> None:
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    rows = [(x["Kernel Name"], float(x["Metric Value"]), x["Grid Size"], x["Block Size"]) for x in csv.DictReader(lines)]
    fw = [i for i, r in enumerate(rows) if "tbe_forward_kernel" in r[0]]
    seg = rows[fw[step_index]:fw[step_index + 1]]
    tot = sum(r[1] for r in seg)
    agg = collections.OrderedDict()
    for n, t, g, b in seg:
        short = re.sub(r"^void ", "", n)
        short = re.sub(r"[<(].*", "", short)[:64]
        a = agg.setdefault(short, [0, 0.0])
        a[0] += 1
        a[1] += t
    print(f"launches in the whole capture: {len(rows)}; steps found: {len(fw) - 1}; step {step_index}: "
          f"{len(seg)} launches, {tot / 1e3:.1f} us serialised (cold cache, per-kernel replay)\n")
    print("| kernel | launches/step | us/step | share |")
    print("|---|---:|---:|---:|")
    for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{n}` | {c} | {t / 1e3:.1f} | {100 * t / tot:.1f}% |")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 3)
