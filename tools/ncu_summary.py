#!/usr/bin/env python
"""Key metrics per kernel from an .ncu-rep (read with `ncu -i ... --page raw --csv`)."""
import csv
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "time_us", 1e-3),
    ("dram__bytes_read.sum", "dram_rd", 1.0),
    ("dram__bytes_write.sum", "dram_wr", 1.0),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_%", 1.0),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2_%", 1.0),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex_%", 1.0),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_%", 1.0),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_%", 1.0),
    ("sm__inst_executed_pipe_tensor.sum", "tensor_inst", 1.0),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ_%", 1.0),
    ("launch__registers_per_thread", "regs", 1.0),
    ("launch__grid_size", "grid", 1.0),
    ("launch__block_size", "block", 1.0),
]


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    print("| kernel | " + " | ".join(k[1] for k in KEYS) + " |")
    print("|---|" + "---|" * len(KEYS))
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")].split("(")[0][-48:]
        cells = []
        for key, _label, scale in KEYS:
            if key in hdr and r[hdr.index(key)] not in ("", "n/a"):
                v = r[hdr.index(key)].replace(",", "")
                u = units[hdr.index(key)]
                try:
                    if _label == "time_us":
                        scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u, 1.0)
                    f = float(v) * scale
                    cells.append(("%.1f" % f) + (" " + u if _label.startswith("dram_") and not _label.endswith("%") else ""))
                except ValueError:
                    cells.append(v)
            else:
                cells.append("-")
        print("| %s | %s |" % (name, " | ".join(cells)))


if __name__ == "__main__":
    main()
