"""Per-source-line executed warp instructions from `ncu --page source --csv --print-source cuda,sass` (first function only)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = None
out = []
fname = ""
for r in rows:
    if r and r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r and r[0] == "Line No":
        hdr = r
        iex, ismp = hdr.index("Instructions Executed"), hdr.index("# Samples")
        continue
    if hdr is None or len(r) <= iex or not r[0].isdigit() or not r[iex].isdigit():
        continue
    out.append((int(r[iex]), (int(r[ismp]) if r[ismp].isdigit() else 0), fname, int(r[0]), r[1].strip()[:110]))
agg = {}
for n, sm, f, ln, src in out:
    a = agg.setdefault((f, ln), [0, 0, src])
    a[0] += n
    a[1] += sm
out = [(v[0], v[1], k[0], k[1], v[2]) for k, v in agg.items()]
if len(sys.argv) > 3:  # by line number
    out.sort(key=lambda o: (o[2], o[3]))
    tot = sum(o[0] for o in out)
    tots = sum(o[1] for o in out)
    for n, sm, f, ln, src in out:
        if 100.0 * n / tot >= float(sys.argv[3]):
            print(f"{100.0*n/tot:5.2f}% i {100.0*sm/max(tots,1):5.2f}% s  {f}:{ln}  {src}")
    sys.exit(0)
tot = sum(o[0] for o in out)
tots = sum(o[1] for o in out)
print("total warp instr", tot, "samples", tots)
for n, s, f, ln, src in sorted(out, reverse=True)[:top]:
    print(f"{100.0*n/tot:5.2f}% i {100.0*s/max(tots,1):5.2f}% s  {f}:{ln}  {src}")
