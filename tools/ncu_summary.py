#!/usr/bin/env python3
"""`ncu -i REP --page raw --csv` -> a compact per-launch table (CSV on stdout) with the metrics the roofline discussion
uses: duration, DRAM bytes and throughput, achieved occupancy, issue activity, registers, L2 hit rate, tensor pipe."""
import csv
import subprocess
import sys

WANT = [("Kernel Name", "kernel"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("launch__registers_per_thread", "regs"), ("gpu__time_duration.sum", "us"),
        ("dram__bytes_read.sum", "dram_read_MB"), ("dram__bytes_write.sum", "dram_write_MB"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
        ("lts__t_sector_hit_rate.pct", "l2_hit_pct"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy_pct"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_active_pct"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_pipe_pct"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct")]


def main(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    cols = [(h, n) for h, n in WANT if h in idx]
    w = csv.writer(sys.stdout)
    w.writerow([n for _, n in cols])
    for r in rows[2:]:
        out = []
        for h, n in cols:
            v = r[idx[h]]
            if n == "kernel":
                v = v.split("(")[0].replace("void ", "").replace("b200::", "")[:60]
            elif n == "us":
                v = "%.2f" % (float(v.replace(",", "")) * {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3,
                                                           "msecond": 1e3}.get(units[idx[h]], 1.0))
            elif n.endswith("_MB"):
                v = "%.2f" % (float(v.replace(",", "")) * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(units[idx[h]], 1.0))
            else:
                try:
                    v = "%.2f" % float(v.replace(",", ""))
                except ValueError:
                    pass
            out.append(v)
        w.writerow(out)


if __name__ == "__main__":
    main(sys.argv[1])
