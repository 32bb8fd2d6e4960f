"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name.
usage: python tools/summarize_launches.py gpurun_out/launches.csv > profiles/xxx_summary.txt"""
import collections
import csv
import re
import sys


def main(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    tot = 0.0
    for row in csv.DictReader(lines):
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(row["Metric Unit"], 1e-6)
        name = re.sub(r"\(.*", "", re.sub(r"<.*", "", row["Kernel Name"]))[:60]
        agg[name][0] += 1
        agg[name][1] += v
        tot += v
    print("# %s : %d launches, %.2f ms total (serialised, cold cache)" % (path, sum(a[0] for a in agg.values()), tot))
    print("# %10s %7s %7s  kernel" % ("ms", "share", "count"))
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%12.3f %6.1f%% %7d  %s" % (t, 100 * t / tot, n, k))


if __name__ == "__main__":
    main(sys.argv[1])
