"""Aggregate an `ncu --page source --csv` export: per kernel, executed warp instructions by opcode and the
instructions with the most stall samples.   python tools/ncu_source.py file.csv [kernel-index] [top-n]"""
import csv
import re
import sys
from collections import Counter


def kernels(path):
    cur, hdr, out = None, None, []
    for row in csv.reader(open(path)):
        if row and row[0] == "Kernel Name":
            cur = {"name": row[1], "rows": []}
            out.append(cur)
            hdr = None
        elif cur is not None and hdr is None and row and row[0] == "Address":
            hdr = row
            cur["hdr"] = hdr
        elif cur is not None and hdr is not None and row:
            cur["rows"].append(row)
    return out


def opcode(s):
    s = re.sub(r"^@!?U?P\d+\s+", "", s.strip())
    return s.split()[0].rstrip(";") if s else "?"


def main():
    ks = kernels(sys.argv[1])
    which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    for i, k in enumerate(ks):
        print(i, k["name"][:100], len(k["rows"]), "instructions")
    k = ks[which]
    h = k["hdr"]
    isrc, iex, ismp = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
    ops, smp = Counter(), Counter()
    tot = 0
    for r in k["rows"]:
        n = int(float(r[iex] or 0))
        tot += n
        base = opcode(r[isrc])
        ops[base.split(".")[0] + ("." + base.split(".")[1] if base.startswith(("IMAD", "LDS", "STS", "LDG", "STG")) and "." in base else "")] += n
        smp[base.split(".")[0]] += int(float(r[ismp] or 0))
    print(f"\nkernel {which}: {tot} warp instructions")
    for o, n in ops.most_common(30):
        print(f"  {o:18s} {n:12d} {100.0 * n / tot:6.2f}%   samples {smp[o.split('.')[0]]}")
    print("\ntop stall-sample instructions:")
    stall_cols = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
    rows = sorted(k["rows"], key=lambda r: -int(float(r[ismp] or 0)))[:top]
    for r in rows:
        st = sorted(((int(float(r[i] or 0)), h[i][6:]) for i in stall_cols), reverse=True)[:3]
        print(f"  {r[0]:>6s} {r[ismp]:>6s} {r[iex]:>9s}  {r[isrc][:70]:70s} {st}")


if __name__ == "__main__":
    main()
