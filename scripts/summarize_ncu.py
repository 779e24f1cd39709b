#!/usr/bin/env python
"""Condense ncu CSV exports into the small tables committed under profiles/.

  python scripts/summarize_ncu.py launches gpurun_out/TAG_launches.csv profiles/TAG_launch_summary.csv
  python scripts/summarize_ncu.py full     gpurun_out/TAG_full_raw.csv profiles/TAG_full_summary.csv
"""
import csv
import re
import sys
from collections import OrderedDict, defaultdict


def short(name):
    name = re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", name)
    name = re.sub(r"^void ", "", name)
    return re.sub(r"\(.*$", "", name)


def rows_of(path):
    with open(path, newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    return list(csv.DictReader(lines))


def launches(src, dst):
    rows = rows_of(src)
    tot, cnt = defaultdict(float), defaultdict(int)
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        us = v / 1e3 if unit in ("ns", "nsecond") else v * (1e3 if unit in ("ms", "msecond") else 1.0)
        k = short(r["Kernel Name"])
        tot[k] += us
        cnt[k] += 1
    total = sum(tot.values())
    with open(dst, "w") as f:
        f.write(f"# {src}: ncu --metrics gpu__time_duration.sum --clock-control none; {sum(cnt.values())} launches, "
                f"{total / 1e3:.1f} ms in kernels (cold-cache, serialised: compare SHARES)\n")
        f.write("kernel,launches,total_us,avg_us,share\n")
        for k in sorted(tot, key=tot.get, reverse=True):
            f.write(f"{k},{cnt[k]},{tot[k]:.1f},{tot[k] / cnt[k]:.1f},{tot[k] / total:.4f}\n")


WANT = OrderedDict([
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram_read"),
    ("dram__bytes_write.sum", "dram_write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
    ("launch__registers_per_thread", "regs"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy_pct"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("lts__t_sector_hit_rate.pct", "l2_hit_pct"),
])


def full(src, dst):
    rows = rows_of(src)
    if not rows:
        sys.exit("empty")
    cols = list(rows[0].keys())
    # --page raw --csv: one row per launch, one column per metric (plus a units row first)
    out = []
    for r in rows:
        if not r.get("ID", "").strip().isdigit():
            continue
        rec = {"id": r["ID"], "kernel": short(r["Kernel Name"])}
        for m, nick in WANT.items():
            c = next((c for c in cols if c == m or c.startswith(m)), None)
            rec[nick] = r.get(c, "") if c else ""
        out.append(rec)
    units = rows[0]
    with open(dst, "w") as f:
        f.write(f"# {src}: ncu --set full --clock-control none, raw page; units: " +
                "; ".join(f"{nick}={units.get(next((c for c in cols if c.startswith(m)), ''), '')}" for m, nick in WANT.items()) + "\n")
        f.write("id,kernel," + ",".join(WANT.values()) + "\n")
        for rec in out:
            f.write(",".join(str(rec[k]).replace(",", "") for k in ["id", "kernel"] + list(WANT.values())) + "\n")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
