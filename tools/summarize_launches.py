#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: total time and launch count per kernel."""
import csv, sys, re, collections

def short(name):
    """Drop template arguments and parameter lists, keep the last two scopes."""
    out, depth = [], 0
    for ch in name:
        if ch in "<(":
            depth += 1
        elif ch in ">)":
            depth -= 1
        elif depth == 0:
            out.append(ch)
    n = "".join(out).replace("void ", "").strip()
    return "::".join(n.split("::")[-2:])[-80:]


def main(path, top=40):
    tot = collections.defaultdict(float); cnt = collections.Counter()
    with open(path, newline="") as f:
        rows = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(rows):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = r["Kernel Name"]
        name = short(name)
        tot[name] += float(r["Metric Value"].replace(",", "")) / 1e3
        cnt[name] += 1
    total = sum(tot.values())
    print(f"total {total:.1f} us over {sum(cnt.values())} launches")
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:top]:
        print(f"{v:10.1f} us {100*v/total:5.1f}% n={cnt[k]:5d} avg={v/cnt[k]:8.2f}  {k}")

if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
