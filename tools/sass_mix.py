"""Instruction mix of one kernel from `ncu --page source --csv --print-source sass` output (warp-level executed counts per opcode)."""
import csv
import re
import sys
from collections import Counter

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ia, isrc, iex, ismp = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
mix, smp = Counter(), Counter()
tot = 0
for r in rows[2:]:
    if r and r[0] == "Kernel Name":
        break  # first launch only
    if len(r) <= iex or not r[iex].isdigit():
        continue
    s = r[isrc].strip()
    s = re.sub(r"^@!?U?P\d+\s+", "", s)
    op = s.split()[0] if s else "?"
    op = ".".join(op.split(".")[:2]) if op.startswith(("VIMNMX", "VIADD", "LDS", "STS", "IMAD", "HFMA2", "HADD2", "HMNMX2")) else op.split(".")[0]
    n = int(r[iex])
    mix[op] += n
    smp[op] += int(r[ismp] or 0)
    tot += n
print("total warp instr", tot)
for op, n in mix.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 30):
    print(f"{op:20s} {n:12d} {100.0*n/tot:6.2f}%  samples {smp[op]}")
