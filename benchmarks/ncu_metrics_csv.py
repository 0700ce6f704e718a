"""Summarise an `ncu --metrics ... --csv --log-file X.csv` launch list: one line per kernel name x grid with the mean of
every collected metric over its launches, plus achieved DRAM GB/s (read + write bytes / duration)."""
import collections
import csv
import sys


def main(path):
    rows = [r for r in csv.reader(open(path)) if len(r) >= 15 and r[0].isdigit()]
    launches = collections.OrderedDict()
    for r in rows:
        launches.setdefault(r[0], {"name": r[4].split("(")[0], "grid": r[8], "block": r[7]})[r[12]] = float(r[14].replace(",", ""))
        launches[r[0]]["unit:" + r[12]] = r[13]
    groups = collections.OrderedDict()
    for l in launches.values():
        groups.setdefault((l["name"], l["grid"], l["block"]), []).append(l)
    for (name, grid, block), ls in groups.items():
        def mean(k):
            return sum(l.get(k, 0.0) for l in ls) / len(ls)
        dur = mean("gpu__time_duration.sum")
        unit = ls[0].get("unit:gpu__time_duration.sum", "ns")
        dur_us = dur / 1e3 if unit in ("ns", "nsecond") else dur * {"us": 1, "usecond": 1, "ms": 1e3, "msecond": 1e3}.get(unit, 1)
        def byt(k):
            u = ls[0].get("unit:" + k, "byte")
            return mean(k) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
        rd, wr = byt("dram__bytes_read.sum"), byt("dram__bytes_write.sum")
        extra = " ".join(f"{k.split('.')[0].split('__')[1]}={mean(k):.1f}" for k in ls[0] if "pct_of_peak" in k and not k.startswith("unit:"))
        print(f"{name[:60]:60s} grid {grid:14s} block {block:12s} n={len(ls):2d} {dur_us:9.1f} us  dram rd {rd / 1e6:8.1f} MB wr {wr / 1e6:8.1f} MB"
              f"  {((rd + wr) / dur_us / 1e3) if dur_us else 0:7.1f} GB/s  {extra}")


if __name__ == "__main__":
    main(sys.argv[1])
