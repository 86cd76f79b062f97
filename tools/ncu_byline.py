"""Samples per SOURCE LINE RANGE of one file from an .ncu-rep (source page, cuda + sass), with the dominant stall reasons.
usage: ncu_byline.py report.ncu-rep file_substring [kernel index]"""
import csv, io, subprocess, sys
rep, want = sys.argv[1], sys.argv[2]; kidx = int(sys.argv[3]) if len(sys.argv) > 3 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
secs = []; i = 0
while i < len(rows):
    if rows[i] and rows[i][0] == "File Path":
        path = rows[i][1]; fn = rows[i+1][1]; hdr = rows[i+2]; j = i + 3; body = []
        while j < len(rows) and not (rows[j] and rows[j][0] == "File Path"):
            body.append(rows[j]); j += 1
        secs.append((path, fn, hdr, body)); i = j
    else:
        i += 1
fns = []
for s in secs:
    if s[1] not in fns: fns.append(s[1])
fn = fns[kidx]
tot = 0; per = {}
for path, f, hdr, body in secs:
    if f != fn: continue
    iS = hdr.index("# Samples"); iE = hdr.index("Instructions Executed")
    stall = [(j, c) for j, c in enumerate(hdr) if c.startswith("stall_") and "Not Issued" not in c]
    for r in body:
        if r[0] == "": continue
        try: s_ = float(r[iS]); e_ = float(r[iE])
        except ValueError: continue
        tot += s_
        if want in path:
            d = per.setdefault(int(r[0]), [0.0, 0.0, {}])
            d[0] += s_; d[1] += e_
            for j, c in stall: d[2][c] = d[2].get(c, 0.0) + float(r[j] or 0)
print("function:", fn[:80], "total samples", tot)
for ln in sorted(per):
    s_, e_, st = per[ln]
    if s_ / tot < 0.002: continue
    top = sorted(st.items(), key=lambda x: -x[1])[:3]
    print(f"{want}:{ln:4d} {s_/tot*100:5.2f}%  exec {e_/1e6:8.1f}M  " + " ".join(f"{c[6:]} {v/max(s_,1)*100:.0f}%" for c, v in top))
