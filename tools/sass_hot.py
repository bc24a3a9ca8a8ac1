"""Hot SASS lines of one kernel from an ncu report's source page.
    python tools/sass_hot.py <report.ncu-rep> <kernel-regex> [min_frac]"""
import csv, subprocess, sys
rep, kre = sys.argv[1], sys.argv[2]
minf = float(sys.argv[3]) if len(sys.argv) > 3 else 0.004
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", f"regex:{kre}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
import os
_his = [i for i, r in enumerate(rows) if "Source" in r and "Address" in r]
hi = _his[int(os.environ.get("KIDX", "0"))]  # KIDX selects the n-th captured launch that matches
hdr = rows[hi]
i_src, i_s, i_ex = hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
data = []
for r in rows[hi + 1:]:
    if len(r) <= i_ex or r[i_ex] == "Instructions Executed" or not r[i_ex].isdigit():
        if len(r) > i_ex and r[0] == "Address":
            break  # next kernel instance
        continue
    data.append((r[i_src].strip(), int(r[i_s] or 0), int(r[i_ex] or 0)))
tot_s = sum(d[1] for d in data) or 1
tot_e = sum(d[2] for d in data) or 1
print("total samples", tot_s, "total warp-inst", tot_e, "sass lines", len(data))
for k, (s, sm, ex) in enumerate(data):
    if ex > minf * tot_e or sm > 0.01 * tot_s:
        print(f"{k:5d} ex={ex/1e6:8.2f}M st={100*sm/tot_s:5.2f}%  {s[:100]}")
