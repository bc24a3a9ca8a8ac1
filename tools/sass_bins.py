"""Stall samples and executed warp-instructions of one kernel, binned over its SASS (ncu source page).
    python tools/sass_bins.py <report.ncu-rep> <kernel-regex> [lines per bin]"""
import csv, subprocess, sys, os
rep, kre = sys.argv[1], sys.argv[2]
nb = int(sys.argv[3]) if len(sys.argv) > 3 else 100
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", f"regex:{kre}"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
his = [i for i, r in enumerate(rows) if "Source" in r and "Address" in r]
hi = his[int(os.environ.get("KIDX", "0"))]
hdr = rows[hi]
i_src, i_s, i_ex = hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
data = []
for r in rows[hi + 1:]:
    if len(r) <= i_ex or not r[i_ex].isdigit():
        if len(r) > i_ex and r[0] == "Address":
            break
        continue
    data.append((r[i_src].strip(), int(r[i_s] or 0), int(r[i_ex] or 0)))
ts = sum(d[1] for d in data) or 1
te = sum(d[2] for d in data) or 1
print("samples", ts, "warp-inst", te, "lines", len(data))
for b in range(0, len(data), nb):
    seg = data[b:b + nb]
    s = sum(d[1] for d in seg); e = sum(d[2] for d in seg)
    top = max(seg, key=lambda d: d[1])
    tags = set()
    for d in seg:
        op = d[0].split()[0] if not d[0].startswith("@") else d[0].split()[1]
        if op.split(".")[0] in ("BAR", "LDG", "STG", "ATOMS", "ATOMG", "REDG", "MATCH", "VOTE", "LDS", "STS", "MUFU", "RED", "ERRBAR", "MEMBAR", "CCTL", "LDL", "STL"):
            tags.add(op.split(".")[0])
    print(f"{b:5d}-{b+len(seg)-1:5d} st={100*s/ts:5.1f}% ex={100*e/te:5.1f}%  top: {top[0][:46]:46s} {','.join(sorted(tags))}")
