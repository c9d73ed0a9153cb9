"""Per-source-line stall samples and executed instructions from `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass`.
usage: python tools/ncu_lines.py file.csv [top]"""
import collections
import csv
import sys


def num(s):
    try:
        return int(s.replace(",", ""))
    except ValueError:
        return 0


rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur, agg = None, collections.OrderedDict()
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1] if len(r) > 1 else None
    elif r[0].isdigit():
        a = agg.setdefault((cur, int(r[0]), r[1][:100]), [0, 0])
        a[0] += num(r[4])
        a[1] += num(r[7])
tot, ti = sum(a[0] for a in agg.values()), sum(a[1] for a in agg.values())
print("total samples", tot, "warp instructions", ti)
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%5.1f%% smp %5.1f%% ins  %s:%d  %s" % (100 * a[0] / max(tot, 1), 100 * a[1] / max(ti, 1), (key[0] or "")[-12:], key[1], key[2]))
