"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel."""
import collections
import csv
import re
import sys


def main(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(row["Metric Unit"], 1.0)
        name = re.sub(r"^void ", "", row["Kernel Name"])
        name = re.sub(r"\(.*", "", name)[:100]
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(t for _, t in agg.values())
    print(f"{'launches':>8} {'avg us':>10} {'total ms':>10} {'share':>7}  kernel")
    for k, (n, t) in agg.items():
        print(f"{n:8d} {t / n / 1e3:10.1f} {t / 1e6:10.3f} {100 * t / tot:6.1f}%  {k}")


if __name__ == "__main__":
    main(sys.argv[1])
