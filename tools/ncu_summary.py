"""Compact per-kernel table from an `ncu --page raw --csv` dump (one row per profiled launch).
python tools/ncu_summary.py raw.csv > profiles/xxx.md"""
import csv
import sys

COLS = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dramR"), ("dram__bytes_write.sum", "dramW"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("l1tex__t_sector_hit_rate.pct", "L1hit%"), ("lts__t_sector_hit_rate.pct", "L2hit%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("smsp__thread_inst_executed_per_inst_executed.ratio", "thr/inst"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"), ("launch__block_size", "block")]
STALLS = ["long_scoreboard", "short_scoreboard", "barrier", "wait", "math_pipe_throttle", "lg_throttle", "mio_throttle",
          "branch_resolving", "no_instruction", "not_selected", "dispatch_stall", "imc_miss", "drain", "membar", "sleeping",
          "tex_throttle", "selected"]


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    print("| # | kernel | " + " | ".join(n for _, n in COLS) + " | top stalls (warps per issue) |")
    print("|---|---|" + "---|" * (len(COLS) + 1))
    for r in rows[2:]:
        name = r[ix["Kernel Name"]].replace("<unnamed>::", "").split("(")[0]
        cells = []
        for m, _ in COLS:
            if m in ix:
                v, u = r[ix[m]], units[ix[m]]
                try:
                    f = float(v.replace(",", ""))
                    v = ("%.3g" % f) if abs(f) < 1e6 else ("%.4g" % f)
                except ValueError:
                    pass
                cells.append(v + (" " + u if u and u not in ("%", "register/thread") and not m.startswith("launch") else ""))
            else:
                cells.append("-")
        st = []
        for s in STALLS:
            m = "smsp__average_warps_issue_stalled_%s_per_issue_active.ratio" % s
            if m in ix:
                try:
                    st.append((float(r[ix[m]]), s))
                except ValueError:
                    pass
        st.sort(reverse=True)
        print("| %s | %s | " % (r[ix["ID"]], name) + " | ".join(cells) + " | " +
              ", ".join("%s %.2f" % (s, v) for v, s in st[:4]) + " |")


if __name__ == "__main__":
    main(sys.argv[1])
