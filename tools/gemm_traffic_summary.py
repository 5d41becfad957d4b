#!/usr/bin/env python
"""Summarises an ncu --csv --page raw capture of the GEMM launches of one training step: per distinct (grid, duration class)
it is enough to report totals - DRAM bytes read/written, time, L2 hit rate - plus the largest launches."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1], errors="replace")))
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
cols = {n: i for i, n in enumerate(rows[hdr])}
units = rows[hdr + 1]
data = [r for r in rows[hdr + 2:] if len(r) == len(rows[hdr])]


def col(name):
    i = cols[name]
    u = units[i]
    scale = {"Gbyte": 1e3, "Mbyte": 1.0, "Kbyte": 1e-3, "byte": 1e-6, "ms": 1e3, "us": 1.0, "ns": 1e-3, "%": 1.0, "s": 1e6}.get(u, 1.0)
    return [float(r[i].replace(",", "")) * scale for r in data]


rd, wr, t = col("dram__bytes_read.sum"), col("dram__bytes_write.sum"), col("gpu__time_duration.sum")
hit = col("lts__t_sector_hit_rate.pct")
print(f"launches {len(data)}  time {sum(t) / 1e3:.2f} ms  dram read {sum(rd) / 1e3:.2f} GB  write {sum(wr) / 1e3:.2f} GB  "
      f"total {(sum(rd) + sum(wr)) / 1e3:.2f} GB  mean L2 hit {sum(hit) / len(hit):.1f}%")
order = sorted(range(len(data)), key=lambda i: -t[i])[:1]
seen = set()
for i in range(len(data)):
    key = (round(t[i] / 20), round((rd[i] + wr[i]) / 50))
    if key in seen or t[i] < 100:
        continue
    seen.add(key)
    print(f"  launch {i:4d}: {t[i]:7.1f} us  read {rd[i]:7.1f} MB  write {wr[i]:7.1f} MB  L2 hit {hit[i]:5.1f}%")
