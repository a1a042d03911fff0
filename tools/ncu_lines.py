"""Aggregates `ncu --page source --csv --print-source cuda,sass` by source line: share of executed warp instructions,
share of stall samples, average active threads. usage: ncu_lines.py report.ncu-rep kernel_regex [top]"""
import csv, subprocess, sys, io
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + kern],
                     capture_output=True, text=True).stdout
agg, fname, hdr = [], None, None
for r in csv.reader(io.StringIO(out)):
    if len(r) >= 2 and r[0] == "File Path":
        fname = r[1].split("/")[-1]; continue
    if len(r) >= 2 and r[0] == "Line No":
        hdr = r; continue
    if not r or hdr is None or r[0] == "":
        continue
    try:
        ln = int(r[0])
    except ValueError:
        continue
    num = lambda x: int(x) if x.strip().isdigit() else 0
    agg.append((num(r[7]), num(r[4]), num(r[8]), fname, ln, r[1].strip()[:120]))
tot = sum(a[0] for a in agg); ts = sum(a[1] for a in agg)
print("warp instructions", tot, "stall samples", ts)
for a in sorted(agg, reverse=True)[:top]:
    print("%5.1f%% inst %5.1f%% smp  thr %4.1f  %s:%d  %s" % (100 * a[0] / tot, 100 * a[1] / max(ts, 1), a[2] / max(a[0], 1), a[3], a[4], a[5]))
