"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: share of time per kernel name.
usage: python benchmarks/launch_summary.py launches.csv [skip_first_n] [take_n]"""
import csv
import collections
import re
import sys


def main(path, skip=0, take=None):
    rows = []
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") == "gpu__time_duration.sum":
            rows.append((r["Kernel Name"], float(r["Metric Value"].replace(",", "")), r["Grid Size"], r["Block Size"]))
    rows = rows[skip:skip + take if take else None]
    agg = collections.OrderedDict()
    for name, ns, grid, block in rows:
        short = re.sub(r"\(.*$", "", name)
        short = re.sub(r"<unnamed>::|void ", "", short)[:90]
        e = agg.setdefault(short, [0.0, 0, set()])
        e[0] += ns
        e[1] += 1
        e[2].add(grid)
    tot = sum(e[0] for e in agg.values())
    print(f"# {len(rows)} launches, total {tot / 1e6:.3f} ms (serialised, cold-cache ncu replay times)")
    for k, e in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print(f"{100 * e[0] / tot:6.2f}%  n={e[1]:5d} avg_us={e[0] / e[1] / 1e3:9.2f}  {k}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 0, int(sys.argv[3]) if len(sys.argv) > 3 else None)
