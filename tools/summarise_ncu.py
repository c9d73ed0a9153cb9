#!/usr/bin/env python
"""Turns the raw ncu outputs of tools/profile_round.sh into the small tracked summaries under profiles/.
  python tools/summarise_ncu.py launches <launches.csv> <out.csv> "<header comment>"
  python tools/summarise_ncu.py full <report.ncu-rep> <out.csv> "<header comment>"
"""
import collections
import csv
import io
import subprocess
import sys

KEEP = ["Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "gpu__time_duration.sum", "launch__block_size",
        "launch__grid_size", "launch__registers_per_thread", "lts__t_sector_hit_rate.pct",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active"]


def launches(src, dst, comment):
    rows = [r for r in csv.reader(l for l in open(src) if not l.startswith("==")) if len(r) > 14]
    hdr = rows[0]
    ik, iv = hdr.index("Kernel Name"), hdr.index("Metric Value")
    tot, cnt = collections.Counter(), collections.Counter()
    for r in rows[1:]:
        tot[r[ik]] += float(r[iv].replace(",", "")) / 1e3
        cnt[r[ik]] += 1
    total = sum(tot.values())
    with open(dst, "w") as f:
        f.write("# " + comment + "\n")
        w = csv.writer(f)
        w.writerow(["kernel", "launches", "total_us", "share_pct", "avg_us"])
        for k, t in tot.most_common():
            w.writerow([k, cnt[k], f"{t:.1f}", f"{100 * t / total:.2f}", f"{t / cnt[k]:.2f}"])


def full(src, dst, comment):
    out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[0]
    idx = [hdr.index(k) for k in KEEP if k in hdr]
    with open(dst, "w") as f:
        f.write("# " + comment + "\n")
        w = csv.writer(f)
        for r in rows:
            w.writerow([r[i] for i in idx])


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](*sys.argv[2:5])
