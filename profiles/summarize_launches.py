"""Launch list of one command (ncu --metrics gpu__time_duration.sum --clock-control none --csv) by kernel:
count, total and mean duration, share of the GPU time of the run.  Per-launch times under ncu are
cold-cache and serialised: the SHARES are what is comparable with a plain run.
Usage: python profiles/summarize_launches.py LAUNCHES.csv > profiles/NAME_launches_summary.txt"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0].isdigit()]
by = collections.OrderedDict()
for r in rows:
    name = re.sub(r"\(.*", "", r[4]).replace("<unnamed>::", "")
    ns = float(r[-1].replace(",", ""))
    c = by.setdefault(name, [0, 0.0, r[7], r[8]])
    c[0] += 1
    c[1] += ns
tot = sum(c[1] for c in by.values())
print(f"{len(rows)} launches, {tot / 1e6:.1f} ms of GPU time under ncu")
print(f"{'kernel':32s} {'launches':>9s} {'total ms':>10s} {'mean us':>10s} {'share':>7s}  block / grid of the first launch")
for name, (n, ns, blk, grd) in sorted(by.items(), key=lambda kv: -kv[1][1]):
    print(f"{name:32s} {n:9d} {ns / 1e6:10.2f} {ns / n / 1e3:10.1f} {100 * ns / tot:6.1f}%  {blk} / {grd}")
