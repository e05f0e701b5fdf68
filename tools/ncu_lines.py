"""Per-source-line and per-phase stall breakdown of window_lm_kernel from an `ncu --set full --import-source on` report.
usage: python tools/ncu_lines.py <report.ncu-rep> [top_n]   (reads the report, runs nothing on a GPU)"""
import csv
import re
import subprocess
import sys
from collections import defaultdict


def ncu(*a):
    return subprocess.run(["ncu", *a], capture_output=True, text=True, check=True).stdout


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    rows = list(csv.reader(ncu("-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass").splitlines()))
    cur = hdr = line = None
    by_line = defaultdict(lambda: defaultdict(float))
    stall_cols = None
    for r in rows:
        if len(r) == 2 and r[0] == "File Path":
            cur = r[1].split("/")[-1]
            continue
        if len(r) > 5 and r[0] == "Line No":
            hdr = r
            stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
            continue
        if hdr and len(r) == len(hdr):
            if r[0] != "":
                line = int(r[0])
                continue
            key = (cur, line)
            d = by_line[key]
            try:
                d["samples"] += int(r[hdr.index("# Samples")])
                d["winst"] += int(r[hdr.index("Instructions Executed")])
            except ValueError:
                pass
            for i in stall_cols:
                try:
                    d[hdr[i]] += int(r[i])
                except ValueError:
                    pass
    tot = sum(d["samples"] for d in by_line.values())
    print("total samples", tot)
    agg = defaultdict(float)
    for d in by_line.values():
        for k, v in d.items():
            if k.startswith("stall_"):
                agg[k] += v
    print("stall mix:", {k[6:]: round(100 * v / tot, 1) for k, v in sorted(agg.items(), key=lambda x: -x[1]) if v > 0.005 * tot})
    print(f"{'file:line':28s} {'smp%':>6s} {'winst':>10s}  top stalls")
    for key, d in sorted(by_line.items(), key=lambda x: -x[1]["samples"])[:top]:
        st = sorted(((k[6:], v) for k, v in d.items() if k.startswith("stall_")), key=lambda x: -x[1])[:4]
        print(f"{key[0]}:{key[1]:<6d}".ljust(28), f"{100 * d['samples'] / tot:6.2f} {int(d['winst']):10d} ",
              " ".join(f"{k}={100 * v / max(d['samples'], 1):.0f}%" for k, v in st))


if __name__ == "__main__":
    main()
