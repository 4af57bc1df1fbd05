#!/usr/bin/env python
"""Per-launch table of an `ncu --set full` report (run here, no GPU needed):

    python tools/ncu_kernel_table.py gpurun_out/x.ncu-rep profiles/r2_x_ncu_kernels.csv

Writes one CSV line per profiled launch: kernel name, gpu__time_duration (us), dram__bytes_read.sum (MB),
dram__bytes_write.sum (MB), gpu__dram_throughput (% of peak), sm__pipe_tensor_cycles_active (%), grid, block,
registers.  bench.py reads dram read + write of the largest launch of a kernel as `traffic`."""
import csv
import subprocess
import sys

UNIT = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}
COLS = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "MB"), ("dram__bytes_write.sum", "MB"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "%"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "%"),
        ("launch__grid_size", ""), ("launch__block_size", ""), ("launch__registers_per_thread", "")]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    with open(out, "w") as f:
        f.write("# kernel, " + ", ".join(f"{c} [{u}]" if u else c for c, u in COLS) + "\n")
        for r in rows[2:]:
            name = r[idx["Kernel Name"]].replace(",", ";")
            name = name.split("(")[0].replace("void ", "").replace("<unnamed>::", "").replace("unnamed>::", "").strip()
            vals = []
            for c, _ in COLS:
                if c not in idx:
                    vals.append("")
                    continue
                v, u = r[idx[c]], units[idx[c]]
                try:
                    x = float(v.replace(",", ""))
                    x *= UNIT.get(u, 1.0) if c.startswith(("dram__bytes", "gpu__time")) else 1.0
                    vals.append(f"{x:.3f}".rstrip("0").rstrip("."))
                except ValueError:
                    vals.append(v)
            f.write(name + ", " + ", ".join(vals) + "\n")
    print(f"{len(rows) - 2} launches -> {out}")


if __name__ == "__main__":
    main()
