"""Summarise `ncu -i X.ncu-rep --page raw --csv` exports: one markdown row per captured launch."""
import csv
import sys

COLS = [
    ("gpu__time_duration.sum", "dur_us", "time"),
    ("launch__grid_size", "grid", None),
    ("launch__block_size", "block", None),
    ("launch__registers_per_thread", "regs", None),
    ("smsp__inst_executed.sum", "warp_inst_M", "mega"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_pct_active", None),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps_active_pct", None),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "threads_per_inst", None),
    ("dram__bytes_read.sum", "dram_rd_MB", "bytes"),
    ("dram__bytes_write.sum", "dram_wr_MB", "bytes"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct", None),
    ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall_short_sb", None),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall_long_sb", None),
    ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall_wait", None),
    ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stall_barrier", None),
]
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
TSCALE = {"ns": 1e-3, "nsecond": 1e-3, "us": 1, "usecond": 1, "ms": 1e3, "msecond": 1e3, "s": 1e6, "second": 1e6}


def num(s):
    try:
        return float(s.replace(",", ""))
    except ValueError:
        return None


def main(path):
    rows = list(csv.reader(open(path)))
    hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    H, U = rows[hdr], rows[hdr + 1]
    idx = {h: i for i, h in enumerate(H)}
    for key, _, _ in COLS:   # some metrics are exported with a section prefix
        if key not in idx:
            m = [h for h in H if h.endswith("." + key)]
            if m:
                idx[key] = idx[m[0]]
    print("| # | kernel | " + " | ".join(c[1] for c in COLS) + " |")
    print("|---|---|" + "---|" * len(COLS))
    for n, r in enumerate(rows[hdr + 2:]):
        if len(r) < len(H):
            continue
        out = []
        for key, _, conv in COLS:
            if key not in idx:
                out.append("-")
                continue
            v, u = num(r[idx[key]]), U[idx[key]]
            if v is None:
                out.append("-")
            elif conv == "bytes":
                out.append("%.2f" % (v * SCALE.get(u, 1) / 1e6))
            elif conv == "mega":
                out.append("%.2f" % (v / 1e6))
            elif conv == "time":
                out.append("%.1f" % (v * TSCALE.get(u, 1)))
            else:
                out.append(("%.0f" % v) if v >= 1000 else ("%.1f" % v))
        print(f"| {n} | {r[idx['Kernel Name']].split('(')[0][-40:]} | " + " | ".join(out) + " |")


if __name__ == "__main__":
    for p in sys.argv[1:]:
        print(f"\n### {p}\n")
        main(p)
