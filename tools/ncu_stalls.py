"""Per-region stall-reason breakdown from an .ncu-rep (source page, SASS): regions are split at barriers / waits like
ncu_summary.py; for each region the share of samples per stall reason.  usage: ncu_stalls.py report.ncu-rep [kernel index]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; kidx = int(sys.argv[2]) if len(sys.argv) > 2 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
secs = []; cur = None
for r in rows:
    if r and r[0] == "Kernel Name": cur = []; secs.append((r[1], cur)); continue
    if cur is not None and r: cur.append(r)
name, sec = secs[kidx]
h = sec[0]; data = sec[1:]
iS = h.index("# Samples"); iSrc = h.index("Source"); iEx = h.index("Instructions Executed")
stall_cols = [(i, c) for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
tot = sum(float(d[iS]) for d in data) or 1
print(name[:70], "samples", tot)
acc = 0; exe = 0; start = 0; st = {c: 0.0 for _, c in stall_cols}
for i, d in enumerate(data):
    acc += float(d[iS]); exe += float(d[iEx])
    for j, c in stall_cols:
        st[c] += float(d[j] or 0)
    src = d[iSrc]
    if any(t in src for t in ('BAR.SYNC', 'SYNCS.PHASECHK', 'WARPSYNC', 'EXIT')) or i == len(data) - 1:
        if acc / tot > 0.01:
            top = sorted(st.items(), key=lambda x: -x[1])[:5]
            print(f"[{start:5d}-{i:5d}] {acc/tot*100:5.1f}% exec {exe/1e6:8.1f}M  " + "  ".join(f"{c[6:]} {v/max(acc,1)*100:.0f}%" for c, v in top))
        acc = 0; exe = 0; start = i + 1; st = {c: 0.0 for _, c in stall_cols}
