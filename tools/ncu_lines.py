"""Per-source-line summary of an ncu report (needs -lineinfo and --import-source on).
usage: python tools/ncu_lines.py report.ncu-rep [file-filter] [top]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; filt = sys.argv[2] if len(sys.argv) > 2 else None; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, vals = rows[0], rows[2]
for key in ["gpu__time_duration.sum", "launch__registers_per_thread", "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__inst_executed.sum"]:
    if key in hdr: print(key, vals[hdr.index(key)])
for i, h in enumerate(hdr):
    if "issue_stalled" in h and h.endswith("per_issue_active.ratio") and float(vals[i] or 0) > 0.2: print("  ", h.split("issue_stalled_")[1].split("_per_")[0], vals[i])
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
cur = None; h2 = None; lines = []
for r in csv.reader(io.StringIO(src)):
    if len(r) == 2 and r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if len(r) == 2: continue
    if r and r[0] == "Line No": h2 = r; continue
    if r and r[0].isdigit() and r[h2.index("# Samples")].isdigit() and r[h2.index("Instructions Executed")].isdigit(): lines.append((cur, int(r[0]), r))
isamp, iex = h2.index("# Samples"), h2.index("Instructions Executed")
tot = sum(int(r[iex]) for f, l, r in lines); ts = sum(int(r[isamp]) for f, l, r in lines)
print("warp instructions", tot, "samples", ts)
byf = {}
for f, l, r in lines:
    a = byf.setdefault(f, [0, 0]); a[0] += int(r[iex]); a[1] += int(r[isamp])
for k, v in sorted(byf.items(), key=lambda kv: -kv[1][0])[:8]: print("  %-28s inst %5.1f%%  samples %5.1f%%" % (k, 100 * v[0] / tot, 100 * v[1] / ts))
sel = [x for x in lines if not filt or x[0] == filt]
for f, l, r in sorted(sel, key=lambda t: -int(t[2][isamp]))[:top]:
    print("%5.2f %5.2f %s:%d  %s" % (100 * int(r[iex]) / tot, 100 * int(r[isamp]) / ts, f, l, r[1].strip()[:100]))
