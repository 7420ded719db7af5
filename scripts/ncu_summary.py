"""Markdown table from an ncu report: one row per kernel (averaged over its captured launches).

    python scripts/ncu_summary.py gpurun_out/prof_all.ncu-rep [peak_GBps] > profiles/rNN_ncu_all_kernels.md
"""
import csv
import io
import subprocess
import sys
from collections import OrderedDict

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


def to_unit(value, unit, want):
    value = float(value.replace(",", "")) if value else 0.0
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3,
             "second": 1e6}.get(unit, 1.0)
    if want.startswith("MB"):
        return value * scale / 1e6
    if want == "us":
        return value * scale
    return value


def main():
    report = sys.argv[1]
    peak = float(sys.argv[2]) if len(sys.argv) > 2 else 6538.9
    out = subprocess.run(["ncu", "-i", report, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    header, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(header)}
    name_col = ix["Kernel Name"]
    acc = OrderedDict()
    for r in rows[2:]:
        if len(r) <= name_col:
            continue
        name = r[name_col]
        short = name.replace("pdlp_b200::", "").replace("<unnamed>::", "").replace("void ", "")
        short = short.split("(")[0]
        entry = acc.setdefault(short, {"count": 0, "sums": [0.0] * len(COLUMNS)})
        entry["count"] += 1
        for c, (metric, want) in enumerate(COLUMNS):
            if metric in ix:
                entry["sums"][c] += to_unit(r[ix[metric]], units[ix[metric]], want)
    print("| kernel | launches | " + " | ".join(w for _, w in COLUMNS) + " | achieved GB/s | % of measured peak |")
    print("|---|---|" + "---|" * (len(COLUMNS) + 2))
    for name, e in acc.items():
        avg = [s / e["count"] for s in e["sums"]]
        us, rd, wr = avg[0], avg[1], avg[2]
        gbs = (rd + wr) * 1e6 / (us * 1e-6) / 1e9 if us > 0 else 0.0
        cells = [f"{avg[0]:.1f}", f"{avg[1]:.1f}", f"{avg[2]:.1f}", f"{avg[3]:.1f}", f"{avg[4]:.1f}", f"{avg[5]:.1f}", f"{avg[6]:.1f}", f"{avg[7]:.0f}", f"{avg[8]:.0f}"]
        print(f"| `{name}` | {e['count']} | " + " | ".join(cells) + f" | {gbs:.0f} | {100 * gbs / peak:.1f} |")


if __name__ == "__main__":
    main()
