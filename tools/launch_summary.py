"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals and shares.
usage: python tools/launch_summary.py launches.csv [first_n_rows]"""
import collections
import csv
import re
import sys


def main():
    rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0].isdigit()]
    if len(sys.argv) > 2:
        rows = rows[:int(sys.argv[2])]
    agg = collections.OrderedDict()
    for r in rows:
        k = re.sub(r"\(.*", "", r[4])
        a = agg.setdefault(k, [0, 0.0, r[7], r[8]])
        a[0] += 1
        a[1] += float(r[14])
    tot = sum(a[1] for a in agg.values())
    print(f"{len(rows)} launches, {tot / 1e6:.3f} ms total (cold-cache, serialised: compare shares, not absolutes)")
    print("| kernel | launches | total ms | share | mean us | block | grid |\n|---|---|---|---|---|---|---|")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k}` | {a[0]} | {a[1] / 1e6:.3f} | {a[1] / tot:.1%} | {a[1] / a[0] / 1e3:.1f} | {a[2]} | {a[3]} |")


if __name__ == "__main__":
    main()
