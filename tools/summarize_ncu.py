"""Key metrics of every kernel in an .ncu-rep (`ncu --set full`), as a markdown table per kernel."""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_tensor.sum"]


def main(path: str) -> None:
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        print(f"### `{r[idx['Kernel Name']][:110]}`\n")
        print("| metric | value | unit |\n|---|---:|---|")
        for w in WANT:
            if w in idx:
                print(f"| {w} | {r[idx[w]]} | {units[idx[w]]} |")
        st = []
        for h, i in idx.items():
            if "issue_stalled" in h and h.endswith("_per_warp_active.pct"):
                try:
                    st.append((float(r[i].replace(",", "")), h))
                except ValueError:
                    pass
        for v, h in sorted(st, reverse=True)[:6]:
            name = h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_warp_active.pct", "")
            print(f"| stall: {name} | {v:.1f} | % of active warps |")
        print()


def traffic_json(path: str) -> None:
    """{kernel base name: mean dram read + write bytes per launch} as JSON (profiles/r*_traffic.json)."""
    import json
    import re
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    acc = {}
    for r in rows[2:]:
        name = re.sub(r"^void\s+", "", r[idx["Kernel Name"]])
        name = re.sub(r"^b200::", "", name).split("<")[0].split("(")[0]
        tot = 0.0
        for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            tot += float(r[idx[m]].replace(",", "")) * scale.get(units[idx[m]], 1)
        acc.setdefault(name, []).append(tot)
    out = {"source": f"{path}: ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum, mean per launch"}
    out.update({k: int(sum(v) / len(v)) for k, v in acc.items()})
    print(json.dumps(out, indent=2))


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[2] == "--traffic-json":
        traffic_json(sys.argv[1])
    else:
        main(sys.argv[1])
