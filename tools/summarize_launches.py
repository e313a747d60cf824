"""Aggregate an `ncu --csv --metrics gpu__time_duration.sum` log into per-kernel totals."""
import csv
import collections
import re
import sys


def main(path, top=40):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    rd = csv.DictReader(lines)
    tot = collections.defaultdict(lambda: [0, 0.0])
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r["Kernel Name"])[:90]
        val = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        val_us = val / 1e3 if unit in ("ns", "nsecond") else (val if unit in ("us", "usecond") else val * 1e3)
        tot[name][0] += 1
        tot[name][1] += val_us
    total = sum(v[1] for v in tot.values())
    print("total_us {:.1f} kernels {}".format(total, sum(v[0] for v in tot.values())))
    for name, (n, us) in sorted(tot.items(), key=lambda kv: -kv[1][1])[:top]:
        print("{:9.1f} us {:5.1f}% x{:<4d} {}".format(us, 100 * us / total, n, name))


def listing(path):
    """Every launch in order: time, grid, kernel (for finding the launches that scale badly with batch size)."""
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        val = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        us = val / 1e3 if unit in ("ns", "nsecond") else (val if unit in ("us", "usecond") else val * 1e3)
        print("{:4s} {:8.1f} us  grid {:>16s}  {}".format(r["ID"], us, r.get("Grid Size", "").replace(" ", ""),
                                                       re.sub(r"\(.*", "", r["Kernel Name"])[:80]))


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[2] == "list":
        listing(sys.argv[1])
    else:
        main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
