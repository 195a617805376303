"""Summarise an ncu launch list of `scripts/bench_configs.py c3|c4` (--metrics gpu__time_duration.sum --csv): the launches
of the last complete step, per kernel.  Usage: python scripts/config_launches.py launches.csv <first-kernel-substring>"""
import collections
import csv
import re
import sys


def main():
    path, first = sys.argv[1], sys.argv[2]
    lines = [l for l in open(path) if not l.startswith("==")]
    rows = list(csv.DictReader(lines))
    names = [(re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "").replace("rtg::", "").replace("<unnamed>::", ""),
              float(r["Metric Value"].replace(",", "")), r["Grid Size"]) for r in rows]
    idx = [i for i, (n, _, _) in enumerate(names) if first in n]
    seg = names[idx[-2]:idx[-1]]
    total = sum(t for _, t, _ in seg)
    print(f"one step: {len(seg)} launches, {total / 1e6:.3f} ms of kernel time (ncu: serialised, cold caches)")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for n, t, _ in seg:
        agg[n][0] += 1
        agg[n][1] += t
    print(f"{'share':>7s} {'us':>10s} {'count':>5s}  kernel")
    for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{100 * t / total:6.1f}% {t / 1e3:10.1f} {c:5d}  {k}")
    print("\nin launch order (first 40):")
    for n, t, g in seg[:40]:
        print(f"{t / 1e3:8.1f} us  {g:14s} {n}")


if __name__ == "__main__":
    main()
