"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv): one steady-state step, per kernel.
Usage: python scripts/launch_summary.py launches.csv [step_index]"""
import collections
import csv
import re
import sys


def main():
    path = sys.argv[1]
    step = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    rows = list(csv.DictReader(lines))
    names = [(r["Kernel Name"], float(r["Metric Value"].replace(",", "")), r["Grid Size"], r["Block Size"]) for r in rows]
    # a step begins after the 256 MiB L2-flush fill kernel that bench.py launches before it
    marks = [i for i, (n, _, _, _) in enumerate(names) if "FillFunctor" in n]
    s, e = marks[step], marks[step + 1]
    seg = names[s + 1:e]
    total = sum(t for _, t, _, _ in seg)
    print(f"step {step}: {len(seg)} launches, {total / 1e6:.3f} ms of kernel time (ncu: serialised, cold caches)")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for n, t, _, _ in seg:
        k = re.sub(r"\(.*", "", n).replace("void ", "").replace("rtg::", "").replace("<unnamed>::", "")
        agg[k][0] += 1
        agg[k][1] += t
    print(f"{'share':>7s} {'us':>10s} {'count':>5s}  kernel")
    for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{100 * t / total:6.1f}% {t / 1e3:10.1f} {c:5d}  {k}")
    g = sum(t for n, t, _, _ in seg if "gemm_tcgen05" in n)
    print(f"gemm_tcgen05_kernel share of the step: {100 * g / total:.1f}%")


if __name__ == "__main__":
    main()
