#!/usr/bin/env python
"""Hot-spot digest of `ncu -i X.ncu-rep --page source --csv` (SASS view): executed warp-instructions
per opcode and per contiguous SASS region, with the sampled stall totals."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
ie = ix["Instructions Executed"]; sm = ix["# Samples"]
tot = sum(int(r[ie] or 0) for r in data); ts = sum(int(r[sm] or 0) for r in data)
print("total warp-instr", tot, "samples", ts, "sass lines", len(data))
ops = collections.Counter(); ops_s = collections.Counter()
for r in data:
    op = r[ix["Source"]].split()[0] if not r[ix["Source"]].startswith("@") else r[ix["Source"]].split()[1]
    op = op.split(".")[0]
    ops[op] += int(r[ie] or 0); ops_s[op] += int(r[sm] or 0)
for op, c in ops.most_common(22):
    print(f"  {op:10s} {c:12d} {c/tot:6.3f}  samples {ops_s[op]/max(ts,1):6.3f}")
chunk = int(sys.argv[2]) if len(sys.argv) > 2 else 400
print("regions:")
for s in range(0, len(data), chunk):
    blk = data[s:s + chunk]
    e = sum(int(r[ie] or 0) for r in blk); sa = sum(int(r[sm] or 0) for r in blk)
    stalls = collections.Counter()
    for r in blk:
        for k in ("stall_barrier", "stall_short_sb", "stall_long_sb", "stall_math", "stall_wait", "stall_mio", "stall_not_selected", "stall_dispatch", "stall_branch_resolving","stall_lg"):
            stalls[k] += int(r[ix[k]] or 0)
    top = ", ".join(f"{k[6:]}:{v/max(sa,1):.2f}" for k, v in stalls.most_common(4))
    print(f"  [{s:6d}] exec {e/tot:6.3f} samples {sa/max(ts,1):6.3f}  {top}   first: {blk[0][ix['Source']][:50]}")
