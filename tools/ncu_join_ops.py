"""Join an ncu summary (tools/ncu_summary.py, kernels in launch order: one ArcFace forward then one generator forward,
tools/ncu_round2.py) with the op tables of tools/profile_ops.py (ops in the same order), so every kernel row carries
the op that launched it.  Ops launch one kernel each, except split-K convs (conv + splitk_reduce_kernel).
usage: python tools/ncu_join_ops.py summary.csv ops.txt > by_op.csv"""
import csv
import re
import sys


def ops_of(path):
    gen, arc, cur = [], [], None
    for line in open(path):
        if line.startswith("== AEI_Net"):
            cur = gen
        elif line.startswith("== iresnet"):
            cur = arc
        else:
            m = re.match(r"\s+(\S+)\s+(conv_tc|aad_tc|conv_simt|stats|upsample|misc)\s+([0-9.]+)", line)
            if m and cur is not None:
                cur.append((m.group(1), m.group(2), float(m.group(3))))
    return arc + gen  # launch order of ncu_round2.py


def main():
    rows = list(csv.DictReader(open(sys.argv[1])))
    ops = ops_of(sys.argv[2])
    out = csv.writer(sys.stdout)
    keep = ["us", "dram_rd_MB", "dram_wr_MB", "dram_pct", "tensor_pct_active", "issue_pct", "l2_hit_pct", "grid"]
    out.writerow(["op", "class", "events_us", "kernel"] + keep)
    i = 0
    for name, cls, ms in ops:
        if i >= len(rows):
            break
        # skip helper kernels that are not ops of their own (pre/post conversions sit outside the op tables)
        want = r"igemm|tapn" if cls in ("conv_tc", "aad_tc") else (r"rgb_const|simt_conv|rgb_conv" if cls == "conv_simt" else None)
        while i < len(rows) and want and not re.search(want, rows[i]["kernel"]):
            out.writerow(["-", "-", "", rows[i]["kernel"]] + [rows[i].get(k, "") for k in keep])
            i += 1
        if i >= len(rows):
            break
        r = rows[i]
        out.writerow([name, cls, "%.1f" % (ms * 1e3), r["kernel"]] + [r.get(k, "") for k in keep])
        i += 1
        if i < len(rows) and "splitk_reduce" in rows[i]["kernel"]:
            r = rows[i]
            out.writerow([name + " (reduce)", cls, "", r["kernel"]] + [r.get(k, "") for k in keep])
            i += 1
    for r in rows[i:]:
        out.writerow(["-", "-", "", r["kernel"]] + [r.get(k, "") for k in keep])


if __name__ == "__main__":
    main()
