"""Executed-instruction mix (by opcode) of one kernel from an ncu report's source page.
    python tools/sass_mix.py <report.ncu-rep> <kernel-regex> [top]"""
import collections, csv, subprocess, sys
rep, kre = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", f"regex:{kre}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
import os
_his = [i for i, r in enumerate(rows) if "Source" in r and "Address" in r]
hi = _his[int(os.environ.get("KIDX", "0"))]  # KIDX selects the n-th captured launch that matches
hdr = rows[hi]
i_src, i_s, i_ex = hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
i_th = hdr.index("Thread Instructions Executed") if "Thread Instructions Executed" in hdr else None
mix = collections.Counter(); st = collections.Counter(); thr = collections.Counter()
for r in rows[hi + 1:]:
    if len(r) <= i_ex or not r[i_ex].isdigit():
        if len(r) > i_ex and r[0] == "Address":
            break
        continue
    src = r[i_src].strip()
    toks = src.split()
    op = toks[1] if toks and toks[0].startswith("@") else (toks[0] if toks else "?")
    op = op.split(".")[0]
    mix[op] += int(r[i_ex]); st[op] += int(r[i_s] or 0)
    if i_th is not None: thr[op] += int(r[i_th] or 0)
tot = sum(mix.values()); ts = sum(st.values()) or 1
print(f"total warp-inst {tot/1e6:.1f}M  samples {ts}")
PIPE = {"FFMA": "fma", "FMUL": "fma", "FADD": "fma", "IMAD": "fma", "FSEL": "alu", "FSETP": "alu", "FMNMX": "alu",
        "SEL": "alu", "LOP3": "alu", "PLOP3": "alu", "ISETP": "alu", "IADD3": "alu", "VIADD": "alu", "MOV": "alu",
        "MUFU": "xu", "SHFL": "lsu", "LDS": "lsu", "STS": "lsu", "LDG": "lsu", "RED": "lsu", "ATOMG": "lsu",
        "R2P": "alu", "P2R": "alu", "VOTE": "alu", "BRA": "cbu", "BSYNC": "cbu", "BSSY": "cbu", "SHF": "alu",
        "LEA": "alu", "FMNMX3": "alu", "VIMNMX": "alu", "I2FP": "alu", "F2I": "xu", "I2F": "xu", "PRMT": "alu"}
pipes = collections.Counter()
for op, c in mix.items(): pipes[PIPE.get(op, "other:" + op)] += c
for op, c in mix.most_common(top):
    print(f"  {op:10s} {c/1e6:9.2f}M {100*c/tot:5.1f}%  stall {100*st[op]/ts:5.1f}%  lanes {thr[op]/c if c and thr[op] else 0:5.1f}")
print({k: f"{100*v/tot:.1f}%" for k, v in pipes.most_common()})
