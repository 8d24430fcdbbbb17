"""Condense `ncu -i X.ncu-rep --page raw --csv` (stdin or a file) into one line per launch with the metrics the
roofline discussion needs.  Usage: ncu -i rep --page raw --csv | python tools/ncu_summary.py > summary.csv"""
import csv
import sys

KEEP = [
    ("gpu__time_duration.sum", "us"),
    ("dram__bytes_read.sum", "dram_rd_MB"),
    ("dram__bytes_write.sum", "dram_wr_MB"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_pct_active"),
    ("sm__inst_executed_pipe_tensor.sum", "tensor_inst"),
    ("sm__pipe_tensor_op_hmma_cycles_active.avg.pct_of_peak_sustained_active", "hmma_pct"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_pct"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps_pct"),
    ("lts__t_sector_hit_rate.pct", "l2_hit_pct"),
    ("lts__t_bytes.sum", "l2_MB"),
    ("l1tex__t_sector_hit_rate.pct", "l1_hit_pct"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("sm__cycles_active.avg", "sm_cycles_active"),
    ("smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct", "stall_long_sb"),
    ("smsp__warp_issue_stalled_barrier_per_warp_active.pct", "stall_barrier"),
    ("smsp__warp_issue_stalled_math_pipe_throttle_per_warp_active.pct", "stall_math"),
    ("smsp__warp_issue_stalled_lg_throttle_per_warp_active.pct", "stall_lg"),
    ("smsp__warp_issue_stalled_mio_throttle_per_warp_active.pct", "stall_mio"),
]


def main():
    f = open(sys.argv[1]) if len(sys.argv) > 1 else sys.stdin
    rows = list(csv.reader(f))
    # header row = first row containing "Kernel Name"; the row after it holds the units
    hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    head, units = rows[hi], rows[hi + 1]
    col = {h: i for i, h in enumerate(head)}
    sel = [(col[m], name, units[col[m]]) for m, name in KEEP if m in col]
    # any tensor-pipe metric the version of ncu exposes
    extra = [(i, h, units[i]) for i, h in enumerate(head) if "pipe_tensor" in h and h not in dict(KEEP) and
             "pct_of_peak_sustained_active" in h]
    out = csv.writer(sys.stdout)
    out.writerow(["id", "kernel"] + [n for _, n, _ in sel] + [h for _, h, _ in extra])
    for r in rows[hi + 2:]:
        if len(r) <= col["Kernel Name"]:
            continue
        vals = []
        for i, n, u in sel + extra:
            v = r[i].replace(",", "")
            try:
                x = float(v)
                if n == "us":
                    x = x / 1e3 if u in ("ns", "nsecond") else (x * 1e3 if u in ("ms", "msecond") else x)
                if n.endswith("_MB"):
                    x = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1e-6) * x
                vals.append("%.4g" % x)
            except ValueError:
                vals.append(v)
        out.writerow([r[col["ID"]], r[col["Kernel Name"]][:60]] + vals)


if __name__ == "__main__":
    main()
