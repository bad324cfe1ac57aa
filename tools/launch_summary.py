"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel name (and optionally grid) totals
over ONE training step (launch window = one stem forward kernel to the next, after the warm-up step)."""
import csv
import re
import sys
from collections import OrderedDict


def load(path):
    rows = []
    with open(path) as fh:
        lines = [l for l in fh if l.startswith('"')]
    for r in csv.DictReader(lines):
        nm = re.sub(r"\(.*", "", r["Kernel Name"]).replace("ubr::", "")
        nm = re.sub(r"^void ", "", nm)
        rows.append((nm, r["Grid Size"], r["Block Size"], float(r["Metric Value"]) / 1e3))
    return rows


def one_step(rows):
    idx = [i for i, r in enumerate(rows) if r[0].startswith(("stem_fwd_kernel", "stem_tc_fwd_kernel"))]
    if len(idx) >= 3:
        return rows[idx[1]:idx[2]]          # the first window is the warm-up step (weight packing, allocations)
    if len(idx) >= 2:
        return rows[idx[0]:idx[1]]
    return rows


def main():
    path = sys.argv[1]
    by_grid = "--grid" in sys.argv
    full = "--list" in sys.argv
    rows = one_step(load(path))
    tot = sum(r[3] for r in rows)
    print("launches in one step: %d, total %.1f us (cold-cache, serialised)" % (len(rows), tot))
    if full:
        for i, r in enumerate(rows):
            print("%4d %-40s grid %-18s block %-12s %10.1f us" % (i, r[0][:40], r[1], r[2], r[3]))
        return
    agg = OrderedDict()
    for nm, grid, blk, us in rows:
        key = (nm, grid) if by_grid else nm
        c, t = agg.get(key, (0, 0.0))
        agg[key] = (c + 1, t + us)
    for key, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-60s n=%4d %10.1f us %5.1f%%" % (str(key)[:60], c, t, 100 * t / tot))


if __name__ == "__main__":
    main()
