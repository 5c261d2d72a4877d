#!/usr/bin/env python
"""Turns the ncu outputs brought back in gpurun_out/ into the small tracked
summaries under profiles/ :

    python scripts/summarize_profiles.py launches <launches.csv> <out.md>
    python scripts/summarize_profiles.py full <prof.ncu-rep> [...]      (prints a table)
"""
import collections
import csv
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram % of peak"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm % of peak"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("launch__registers_per_thread", "registers/thread"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
]


def launches(path, out):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    tot = 0.0
    for row in csv.DictReader(lines):
        val = float(row["Metric Value"].replace(",", ""))
        unit = row["Metric Unit"]
        us = val / 1e3 if unit == "ns" else val * 1e3 if unit == "ms" else val * 1e6 if unit == "s" else val
        name = row["Kernel Name"].split("(")[0].replace("void ", "").replace("<unnamed>::", "")
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += us
        tot += us
    with open(out, "w") as f:
        f.write("| kernel | launches | total us | avg us | share |\n|---|---|---|---|---|\n")
        for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
            f.write("| `%s` | %d | %.1f | %.1f | %.1f %% |\n" % (k, n, t, t / n, 100 * t / tot))
        f.write("\ntotal kernel time %.1f us over %d launches (ncu: cold-cache, serialised; "
                "compare shares, not absolutes)\n" % (tot, sum(a[0] for a in agg.values())))
    print(open(out).read())


def full(paths):
    for p in paths:
        raw = subprocess.run(["ncu", "-i", p, "--page", "raw", "--csv"], stdout=subprocess.PIPE,
                             text=True).stdout
        rows = list(csv.reader(raw.splitlines()))
        hdr, units = rows[0], rows[1]
        for vals in rows[2:]:
            name = vals[hdr.index("Kernel Name")].split("(")[0]
            print("### %s  (%s)" % (name, p))
            for key, label in KEYS:
                if key in hdr:
                    i = hdr.index(key)
                    print("- %s: %s %s" % (label, vals[i], units[i]))
            print()


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3])
    else:
        full(sys.argv[2:])
