#!/usr/bin/env python3
"""Opcode mix of one kernel from `ncu -i rep --page source --csv` output (stdin): executed warp
instructions and stall samples per SASS opcode."""
import csv, sys, collections
rows = list(csv.reader(sys.stdin))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hdr]
ci, si, ss = h.index("Instructions Executed"), h.index("Source"), h.index("# Samples")
ops, smp = collections.Counter(), collections.Counter()
for r in rows[hdr + 1:]:
    if len(r) <= ci or not r[ci].isdigit():
        continue
    toks = r[si].split()
    op = toks[1] if toks and toks[0].startswith("@") else (toks[0] if toks else "?")
    op = ".".join(op.split(".")[:2]) if op.startswith(("IMAD", "IADD3", "LDG", "LDS", "STS", "STG")) else op.split(".")[0]
    ops[op] += int(r[ci]); smp[op] += int(r[ss])
tot, tots = sum(ops.values()), sum(smp.values()) or 1
print("total warp instructions", tot)
for op, n in ops.most_common(25):
    print("  %-14s %12d  %5.1f%%   samples %5.1f%%" % (op, n, 100.0 * n / tot, 100.0 * smp[op] / tots))
