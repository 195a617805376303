"""Summarise an .ncu-rep: duration, DRAM traffic, tensor-pipe activity, stall ranking and the hottest source lines.
Usage: python scripts/ncu_summary.py report.ncu-rep [n_lines]"""
import csv
import io
import subprocess
import sys


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    return rows[0], rows[1], rows[2:]


def main():
    rep = sys.argv[1]
    nlines = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    h, u, vals = raw(rep)
    for v in vals:
        d = dict(zip(h, v))
        print("kernel:", d.get("Kernel Name", "")[:110], "grid", d.get("Grid Size"), "block", d.get("Block Size"))
        for k in ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
                  "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sectors.sum",
                  "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
                  "sm__inst_executed_pipe_tensor_subpipe_hmma.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
                  "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
                  "launch__registers_per_thread", "sm__cycles_active.avg", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]:
            if k in d:
                print(f"  {k:80s} {d[k]:>16s} {u[h.index(k)]}")
        st = {k.replace("smsp__pcsamp_warps_issue_stalled_", ""): float(x) for k, x in d.items()
              if "pcsamp_warps_issue_stalled" in k and "not_issued" not in k and x}
        tot = sum(st.values()) or 1
        print("  stalls:", ", ".join(f"{k} {100 * x / tot:.0f}%" for k, x in sorted(st.items(), key=lambda kv: -kv[1])[:8]))
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    while rows and "Source" not in rows[0]:
        rows = rows[1:]
    if len(rows) < 2:
        return
    hh = rows[0]
    try:
        si = hh.index("# Samples") if "# Samples" in hh else [i for i, c in enumerate(hh) if "Samples" in c][0]
    except Exception:
        print(hh)
        return
    src = hh.index("Source") if "Source" in hh else 1
    body = [r for r in rows[1:] if len(r) > si and r[si].strip().isdigit()]
    total = sum(int(r[si]) for r in body) or 1
    print(f"  hottest lines (of {total} samples):")
    idx = {id(r): i for i, r in enumerate(body)}
    for r in sorted(body, key=lambda r: -int(r[si]))[:nlines]:
        i = idx[id(r)]
        prev = body[i - 1][src].strip()[:60] if i > 0 else ""
        print(f"   {100 * int(r[si]) / total:5.1f}%  #{i:5d} {r[src].strip()[:90]:90s} | prev: {prev}")


if __name__ == "__main__":
    main()
