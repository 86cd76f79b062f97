"""Per-source-line samples and executed instructions from `ncu --page source --print-source cuda,sass`."""
import csv, subprocess, sys, io
rep = sys.argv[1]; kidx = int(sys.argv[2]) if len(sys.argv) > 2 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
# sections: ("File Path", path), ("Function Name", fn), header, lines...
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
print("function:", fn[:90])
lines = []
tot = 0
for path, f, hdr, body in secs:
    if f != fn: continue
    iS = hdr.index("# Samples"); iE = hdr.index("Instructions Executed")
    for r in body:
        if r[0] != "":   # source line row (aggregated)
            try:
                s_, e_ = float(r[iS]), float(r[iE])
            except ValueError:
                continue
            lines.append((s_, e_, path.split('/')[-1], r[0], r[1].strip()[:100]))
            tot += s_
lines.sort(reverse=True)
print("total samples", tot)
for s_, e_, p, ln, src in lines[:45]:
    print(f"{s_/tot*100:5.1f}%  exec {e_/1e6:9.1f}M  {p}:{ln}  {src}")
