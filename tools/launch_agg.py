"""Aggregate an ncu launch list (--metrics gpu__time_duration.sum --csv) between two occurrences of a marker kernel.
usage: python tools/launch_agg.py launches.csv [marker-substring] [--seq N]"""
import collections
import csv
import re
import sys


def main():
    path = sys.argv[1]
    marker = sys.argv[2] if len(sys.argv) > 2 and not sys.argv[2].startswith("--") else None
    nseq = int(sys.argv[sys.argv.index("--seq") + 1]) if "--seq" in sys.argv else 0
    with open(path) as f:
        rows = list(csv.DictReader([l for l in f if not l.startswith("==")]))
    scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}
    seg = rows
    if marker:
        idx = [i for i, r in enumerate(rows) if marker in r["Kernel Name"]]
        seg = rows[idx[0]:idx[1]] if len(idx) > 1 else rows[idx[0]:]
    agg = collections.OrderedDict()
    for r in seg:
        name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("b2::<unnamed>::", "")
        v = float(r["Metric Value"].replace(",", "")) * scale.get(r["Metric Unit"], 1e-3)
        a = agg.setdefault(name, [0, 0.0, 0.0, 1e18])
        a[0] += 1; a[1] += v; a[2] = max(a[2], v); a[3] = min(a[3], v)
    tot = sum(a[1] for a in agg.values())
    print(f"{'kernel':46s} {'n':>5s} {'total us':>10s} {'share':>6s} {'max':>8s} {'min':>8s}")
    for n, a in agg.items():
        print(f"{n[:46]:46s} {a[0]:5d} {a[1]:10.1f} {a[1] / tot:6.3f} {a[2]:8.1f} {a[3]:8.1f}")
    print(f"{'sum':46s} {sum(a[0] for a in agg.values()):5d} {tot:10.1f}")
    for r in seg[:nseq]:
        print(re.sub(r"\(.*", "", r["Kernel Name"])[-40:], r["Grid Size"], r["Metric Value"], r["Metric Unit"])


if __name__ == "__main__":
    main()
