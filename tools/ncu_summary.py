"""Summarise an .ncu-rep: key raw metrics per launch + stall samples aggregated between barriers / by opcode."""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
want = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'lts__t_bytes.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.sum', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_tensor.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__cycles_elapsed.avg',
        'lts__t_sector_hit_rate.pct', 'sm__throughput.avg.pct_of_peak_sustained_elapsed']
want += [h for h in hdr if 'issue_stalled' in h and h.endswith('per_issue_active.ratio')]
for r in rows[2:]:
    print('----')
    for w in want:
        if w in hdr:
            i = hdr.index(w)
            try:
                v = float(r[i]);
                if 'stalled' in w and v < 0.05: continue
            except ValueError:
                pass
            print(f"{w} [{rows[1][i]}] {r[i]}")
sass = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(sass)))
secs = []; cur = None
for r in rows:
    if r and r[0] == "Kernel Name": cur = []; secs.append((r[1], cur)); continue
    if cur is not None and r: cur.append(r)
for name, sec in secs:
    h = sec[0]; data = sec[1:]
    iS = h.index("# Samples"); iSrc = h.index("Source"); iEx = h.index("Instructions Executed")
    tot = sum(float(d[iS]) for d in data) or 1
    print("==== source regions:", name[:60], "samples", tot, "instrs", len(data))
    acc = 0; start = 0; exe = 0
    for i, d in enumerate(data):
        acc += float(d[iS]); exe += float(d[iEx])
        src = d[iSrc]
        if any(t in src for t in ('BAR.SYNC', 'SYNCS.PHASECHK', 'WARPSYNC', 'EXIT')) or i == len(data) - 1:
            if acc / tot > 0.01:
                print(f"  [{start:5d}-{i:5d}] {acc/tot*100:5.1f}%  exec {exe/1e6:9.1f}M  ends: {src.strip()[:60]}")
            acc = 0; exe = 0; start = i + 1
