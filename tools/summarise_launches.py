"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel.

    python tools/summarise_launches.py gpurun_out/launches.csv [steps]
"""
import collections
import csv
import re
import sys


def main(path, steps=None):
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    tot = collections.defaultdict(lambda: [0, 0.0])
    for r in csv.DictReader(lines):
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        try:
            v = float(r["Metric Value"].replace(",", ""))
        except ValueError:
            continue
        if v != v:
            continue
        unit = r["Metric Unit"]
        us = v / 1e3 if unit in ("ns", "nsecond") else (v if unit.startswith("us") else v * 1e3)
        name = re.sub(r"\(.*", "", r["Kernel Name"])[:100]
        tot[name][0] += 1
        tot[name][1] += us
    total = sum(v[1] for v in tot.values())
    ours = sum(v[1] for k, v in tot.items() if "at::" not in k)
    print("| launches | total us | share of all | us / launch | kernel |")
    print("|---:|---:|---:|---:|---|")
    for k, (n, us) in sorted(tot.items(), key=lambda kv: -kv[1][1])[:30]:
        print(f"| {n} | {us:.1f} | {100 * us / total:.1f}% | {us / n:.1f} | `{k}` |")
    print(f"\nTotal {total:.0f} us, of which our kernels {ours:.0f} us "
          f"({sum(v[0] for v in tot.values())} launches).")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else None)
