"""Markdown table of the L1 / DRAM / stall metrics that bound the row-stream (SpMV) kernels, from an
`ncu --page raw --csv` export (one column per captured launch).

    python scripts/ncu_rowstream_md.py profiles/ncu/r01_rowstream_c3_v6.raw.csv > table.md
"""
import csv
import sys

ROWS = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM written"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput, % of peak"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 throughput, % of peak"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "L1 data-pipe wavefronts, % of peak"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "shared-memory wavefronts"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "of which bank conflicts"),
    ("l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum", "global-load wavefronts"),
    ("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "global-load sectors"),
    ("l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "global-load requests"),
    ("smsp__sass_l1tex_data_pipe_lsu_wavefronts_mem_shared_op_ldgsts.sum", "cp.async (LDGSTS) shared wavefronts"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("launch__registers_per_thread", "registers / thread"),
    ("launch__grid_size", "grid"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long scoreboard / issue"),
    ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall short scoreboard / issue"),
    ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stall barrier / issue"),
    ("smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "stall MIO throttle / issue"),
    ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall wait / issue"),
]


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    launches = rows[2:]
    names = []
    for r in launches:
        n = r[ix["Kernel Name"]]
        n = n.replace("pdlp_b200::", "").replace("<unnamed>::", "").replace("void ", "").split("(")[0]
        names.append(n)
    print("| metric | " + " | ".join(f"`{n}`" for n in names) + " |")
    print("|---|" + "---|" * len(names))
    for metric, label in ROWS:
        if metric not in ix:
            continue
        cells = []
        for r in launches:
            v, u = r[ix[metric]], units[ix[metric]]
            try:
                f = float(v.replace(",", ""))
                v = f"{f:,.0f}" if abs(f) >= 1000 else f"{f:.2f}"
            except ValueError:
                pass
            cells.append((v + " " + u).strip())
        print(f"| {label} | " + " | ".join(cells) + " |")


if __name__ == "__main__":
    main()
