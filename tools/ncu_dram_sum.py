#!/usr/bin/env python
"""Sum dram__bytes_read / dram__bytes_write / gpu__time_duration over ALL launches of an
`ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv` log (long format):

    python tools/ncu_dram_sum.py gpurun_out/x.csv [label]
"""
import csv
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0,
        "nsecond": 1e-9, "usecond": 1e-6, "msecond": 1e-3, "second": 1.0}
rows = [l for l in open(sys.argv[1]) if not l.startswith("==")]
rd = csv.DictReader(rows)
tot, launches = {}, set()
for r in rd:
    name, unit, val = r.get("Metric Name"), r.get("Metric Unit"), r.get("Metric Value")
    if not name:
        continue
    launches.add(r.get("ID"))
    tot[name] = tot.get(name, 0.0) + float(val.replace(",", "")) * UNIT.get(unit, 1.0)
label = sys.argv[2] if len(sys.argv) > 2 else sys.argv[1]
print(f"{label}: {len(launches)} launches, dram read {tot.get('dram__bytes_read.sum', 0) / 1e6:.1f} MB, "
      f"dram write {tot.get('dram__bytes_write.sum', 0) / 1e6:.1f} MB, "
      f"sum of kernel durations {tot.get('gpu__time_duration.sum', 0) * 1e3:.3f} ms")
