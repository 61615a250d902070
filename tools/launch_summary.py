"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list:
per kernel name: launches, total ms, share.  usage: launch_summary.py file.csv [first_n_launches]"""
import collections
import csv
import re
import sys


def main():
    path = sys.argv[1]
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    rows = list(csv.DictReader(lines))
    if len(sys.argv) > 2:
        rows = rows[:int(sys.argv[2])]
    agg = collections.OrderedDict()
    tot = 0.0
    for row in rows:
        name = re.sub(r"\(.*", "", row["Kernel Name"])
        name = re.sub(r"^void |\(anonymous namespace\)::|<unnamed>::", "", name)[:80]
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1e-6)
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
        tot += v
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%5d  %9.3f ms  %5.1f%%  %s" % (n, t, 100 * t / tot, k))
    print("total %.3f ms over %d launches" % (tot, len(rows)))


if __name__ == "__main__":
    main()
