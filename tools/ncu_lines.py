"""Summarise an ncu report per CUDA source line: share of executed warp instructions and stall samples."""
import csv, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
hdr = None; agg = {}
for r in rows:
    if r and r[0] == 'Line No': hdr = r; continue
    if hdr is None or len(r) < 8: continue
    if r[2] != '-': continue          # only the per-source-line summary rows
    line = r[0]; src = r[1].strip()
    try: inst = float(r[7] or 0); samp = float(r[6] or 0)
    except ValueError: continue
    a = agg.setdefault((line, src), [0, 0]); a[0] += inst; a[1] += samp
tot = sum(a[0] for a in agg.values()); tots = sum(a[1] for a in agg.values())
print("total warp instructions %.4g, samples %d" % (tot, tots))
for (line, src), (inst, samp) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%5.1f%% inst %5.1f%% samp  L%-4s %s" % (100 * inst / tot, 100 * samp / max(tots, 1), line, src[:105]))
