"""Aggregate an `ncu --page source --csv` SASS dump into contiguous regions of similar execution count.

    ncu -i rep.ncu-rep --page source --csv > src.csv ; python tools/sass_regions.py src.csv [listing.txt]
"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
R = [(int(r[ix['Instructions Executed']]), r[ix['Avg. Threads Executed']][:5], r[ix['# Samples']], r[ix['Source']].strip())
     for r in rows[2:]]
if len(sys.argv) > 2:
    with open(sys.argv[2], "w") as f:
        for n, r in enumerate(R):
            f.write("%4d %9d %5s %5s  %s\n" % ((n,) + r))
tot = sum(r[0] for r in R)
regions, cur = [], 0
for i in range(1, len(R) + 1):
    if i == len(R) or abs(R[i][0] - R[i - 1][0]) > 0.3 * max(R[i][0], R[i - 1][0], 1):
        regions.append((cur, i, sum(r[0] for r in R[cur:i]), sum(int(r[2]) for r in R[cur:i])))
        cur = i
regions.sort(key=lambda x: -x[2])
print("total warp instructions", tot)
for a, b, s, smp in regions[:30]:
    print("%5d-%5d n=%4d sum=%10d (%4.1f%%) samples=%6d cnt~%d thr~%s  %s" % (a, b, b - a, s, 100 * s / tot, smp, R[a][0], R[a][1], R[a][3][:60]))
