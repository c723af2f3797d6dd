"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name.
    python profiles/launch_summary.py gpurun_out/launches.csv profiles/r1_xxx_launches.csv
"""
import csv
import re
import sys
from collections import OrderedDict


def main(src, dst):
    rows = [r for r in csv.reader(open(src, errors="replace")) if len(r) > 14 and r[0].isdigit()]
    agg = OrderedDict()
    for r in rows:
        name, unit, val = r[4], r[13], float(r[14].replace(",", ""))
        us = val * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1e-3)
        short = re.sub(r"\(.*", "", name)[:110]
        a = agg.setdefault(short, [0, 0.0])
        a[0] += 1
        a[1] += us
    total = sum(a[1] for a in agg.values())
    with open(dst, "w", newline="") as fh:
        wr = csv.writer(fh)
        wr.writerow(["kernel", "launches", "total_us", "mean_us", "share_pct"])
        for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            wr.writerow([k, n, "%.1f" % us, "%.2f" % (us / n), "%.2f" % (100 * us / total)])
    print("%d launches, %.1f ms total -> %s" % (len(rows), total / 1e3, dst))


if __name__ == "__main__":
    main(*sys.argv[1:3])
