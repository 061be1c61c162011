"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel name.
Usage: python tools/launch_summary.py gpurun_out/launches.csv "<command>" > profiles/launches_summary.csv"""
import collections
import csv
import sys


def main(path, cmd):
    lines = open(path).read().splitlines()
    start = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
    agg, n = collections.OrderedDict(), 0
    for row in csv.DictReader(lines[start:]):
        if row["Metric Name"] != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1.0, "nsecond": 1.0, "us": 1e3, "usecond": 1e3, "ms": 1e6, "msecond": 1e6, "s": 1e9,
              "second": 1e9}[row["Metric Unit"]]
        name = row["Kernel Name"].split("(")[0]
        name = name.replace("void ", "").replace("<unnamed>::", "")[:100]
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
        n += 1
    tot = sum(a[1] for a in agg.values())
    print("# ncu launch list (cold-cache, serialised: compare SHARES, not absolutes)")
    print(f"# command: {cmd}")
    print(f"# {n} launches captured, total {tot / 1e6:.1f} ms")
    print("kernel,launches,total_ms,share")
    for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"\"{k}\",{c},{t / 1e6:.3f},{t / tot:.4f}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "")
