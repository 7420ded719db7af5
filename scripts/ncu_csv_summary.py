"""Markdown table from an `ncu --page raw --csv` export: one row per captured launch.

    python scripts/ncu_csv_summary.py profiles/ncu/r01_every_kernel_c3_v10.raw.csv [peak_GBps]
"""
import csv
import json
import os
import sys

COLUMNS = [
    ("gpu__time_duration.sum", "us"),
    ("dram__bytes_read.sum", "MB read"),
    ("dram__bytes_write.sum", "MB written"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 %"),
    ("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "L1 wavefronts %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
]
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}


def short(name):
    import re
    name = name.replace("pdlp_b200::", "").replace("<unnamed>::", "").replace("(anonymous namespace)::", "").replace("void ", "")
    name = name.split("(")[0]
    name = re.sub(r"_GLOBAL__N__[0-9a-f_]+_cu_[0-9a-f]+", "", name)
    return name[:90]


def main():
    path = sys.argv[1]
    peak = float(sys.argv[2]) if len(sys.argv) > 2 else None
    if peak is None:
        mp = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
        peak = json.load(open(mp))["hbm_gbs"] if os.path.exists(mp) else 6650.0
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    dem = ix.get("Demangled Name", ix.get("Kernel Name"))
    print("| kernel | " + " | ".join(w for _, w in COLUMNS) + " | DRAM GB/s | % of measured peak |")
    print("|---|" + "---|" * (len(COLUMNS) + 2))
    for r in rows[2:]:
        vals = []
        for metric, want in COLUMNS:
            if metric not in ix:
                vals.append(0.0)
                continue
            v = float(r[ix[metric]].replace(",", "") or 0)
            u = units[ix[metric]]
            if want.startswith("MB"):
                v = v * SCALE.get(u, 1.0) / 1e6
            elif want == "us":
                v = v * SCALE.get(u, 1.0)
            vals.append(v)
        us, rd, wr = vals[0], vals[1], vals[2]
        gbs = (rd + wr) * 1e6 / (us * 1e-6) / 1e9 if us > 0 else 0.0
        cells = [f"{vals[0]:.1f}", f"{vals[1]:.1f}", f"{vals[2]:.1f}", f"{vals[3]:.1f}", f"{vals[4]:.1f}", f"{vals[5]:.1f}", f"{vals[6]:.1f}", f"{vals[7]:.0f}", f"{vals[8]:.0f}"]
        print(f"| `{short(r[dem])}` | " + " | ".join(cells) + f" | {gbs:.0f} | {100 * gbs / peak:.1f} |")


if __name__ == "__main__":
    main()
