"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list by kernel, optionally per
phase (phases are split on marker kernels).  usage: summarize_launches.py launches.csv [--from I --to J]"""
import csv, sys, re, argparse, collections

def load(path):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    rd = csv.DictReader(lines)
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        us = v / 1000 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1000)
        name = re.sub(r"\(.*", "", r["Kernel Name"])
        name = name.replace("mt::", "")
        rows.append((int(r["ID"]), name, us, r.get("Grid Size", ""), r.get("Block Size", "")))
    return rows

def summarize(rows, title):
    agg = collections.OrderedDict()
    for _, n, us, g, b in rows:
        a = agg.setdefault(n, [0, 0.0])
        a[0] += 1; a[1] += us
    tot = sum(a[1] for a in agg.values())
    print(f"## {title}: {len(rows)} launches, {tot:.1f} us serialised")
    print("| kernel | launches | us | us/launch | share |\n|---|---:|---:|---:|---:|")
    for n, (c, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{n}` | {c} | {us:.1f} | {us / c:.2f} | {100 * us / tot:.1f}% |")
    print()

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--lo", type=int, default=0)
    ap.add_argument("--hi", type=int, default=1 << 60)
    ap.add_argument("--list", action="store_true")
    a = ap.parse_args()
    rows = [r for r in load(a.csv) if a.lo <= r[0] < a.hi]
    if a.list:
        for r in rows:
            print(r[0], r[1], f"{r[2]:.2f}", r[3], r[4])
    else:
        summarize(rows, f"{a.csv} [{a.lo},{a.hi})")
