#!/usr/bin/env python
"""Print the hottest SASS instructions (by executed count and by stall samples) of one kernel in an .ncu-rep."""
import csv, subprocess, sys
rep, kernel = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kernel], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
iS, iI, iN = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
body = [r for r in rows[2:] if len(r) > iN and r[iI].isdigit()]
tot = sum(int(r[iI]) for r in body); samp = sum(int(r[iN]) for r in body)
print(f"{kernel}: {len(body)} SASS instrs, {tot} warp-instr executed, {samp} samples")
ops = {}
for r in body:
    op = r[iS].split()[0] if not r[iS].strip().startswith("@") else r[iS].split()[1]
    ops[op] = ops.get(op, 0) + int(r[iI])
print("by opcode:", ", ".join(f"{k}={v*100//max(1,tot)}%" for k, v in sorted(ops.items(), key=lambda x: -x[1])[:14]))
print("-- top by stall samples")
for r in sorted(body, key=lambda r: -int(r[iN]))[:top]:
    print(f"{int(r[iN]):6d} {int(r[iI]):9d}  {r[iS].strip()[:100]}")
