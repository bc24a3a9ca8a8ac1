"""Summarise gpurun_out ncu artefacts into profiles/ (text, committed).

    python tools/ncu_summary.py <tag>      # reads gpurun_out/<tag>_launches.csv, <tag>_prof.ncu-rep
"""
import collections
import csv
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
out = open(os.path.join(ROOT, "profiles", f"{tag}_ncu_summary.txt"), "w")

lc = os.path.join(ROOT, "gpurun_out", f"{tag}_launches.csv")
if os.path.exists(lc):
    lines = [l for l in open(lc) if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", row["Kernel Name"])[:80]
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
        agg[name][0] += 1
        agg[name][1] += v
    tot = sum(v[1] for v in agg.values())
    out.write(f"# launch list ({tag}): ncu --metrics gpu__time_duration.sum --clock-control none, "
              f"python bench.py --steps 2 --warmup 3 --no-cpu-baseline\n")
    out.write(f"# total {tot/1e3:.2f} ms over {sum(v[0] for v in agg.values())} launches "
              f"(cold-cache, serialised: compare shares)\n")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.write(f"{v[1]:10.1f} us {100*v[1]/tot:5.1f}% n={v[0]:4d} avg={v[1]/v[0]:8.1f} us  {k}\n")

for rep in (os.path.join(ROOT, "gpurun_out", f"{tag}_prof.ncu-rep"),
            os.path.join(ROOT, "gpurun_out", f"{tag}_loss.ncu-rep"),
            os.path.join(ROOT, "gpurun_out", f"{tag}_densify.ncu-rep"),
            os.path.join(ROOT, "gpurun_out", f"{tag}_compact.ncu-rep"),
            os.path.join(ROOT, "gpurun_out", f"{tag}_assemble.ncu-rep")):
    if not os.path.exists(rep):
        continue
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
            "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
            "smsp__inst_executed.sum", "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_alu.sum",
            "sm__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_lsu.sum",
            "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
            "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
            "smsp__thread_inst_executed_per_inst_executed.ratio"]
    stall = [h for h in hdr if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("not_issued")]
    out.write(f"\n# ncu --set full --clock-control none ({os.path.basename(rep)}); one block per captured launch\n")
    for r in rows[2:]:
        out.write("-----\n" + r[idx["Kernel Name"]][:90] + "\n")
        for w in want:
            if w in idx:
                out.write(f"  {w:66s} {r[idx[w]]:>18s} {units[idx[w]]}\n")
        try:
            sv = sorted(((float(r[idx[h]].replace(",", "")), h.replace("smsp__pcsamp_warps_issue_stalled_", ""))
                         for h in stall), reverse=True)
            tot = sum(v for v, _ in sv) or 1.0
            out.write("  warp-state samples: " + ", ".join(f"{n} {100*v/tot:.0f}%" for v, n in sv[:7]) + "\n")
        except ValueError:
            pass
out.close()
print(open(os.path.join(ROOT, "profiles", f"{tag}_ncu_summary.txt")).read()[-3000:])
