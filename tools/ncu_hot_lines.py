"""Per-CUDA-source-line shares of executed instructions and stall samples from an ncu report captured with
--import-source on (kernels built with -lineinfo):
    ncu -i rep.ncu-rep --page source --csv --print-source sass,cuda > src.csv ; python tools/ncu_hot_lines.py src.csv [N]
"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur = None
hdr = None
lines = []
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r and r[0] == "Line No":
        hdr = r
        continue
    if hdr and len(r) == len(hdr) and r[2] == "-":      # a CUDA source line (SASS rows carry an address)
        lines.append((cur, r))
iI = hdr.index("Instructions Executed")
iW = hdr.index("Warp Stall Sampling (All Samples)")
tot = sum(int(r[iI] or 0) for _, r in lines)
totw = sum(int(r[iW] or 0) for _, r in lines)
print("total warp-instructions %d, stall samples %d" % (tot, totw))
bf = collections.Counter()
bw = collections.Counter()
for f, r in lines:
    bf[f] += int(r[iI] or 0)
    bw[f] += int(r[iW] or 0)
for f, c in bf.most_common():
    print("%5.1f%% instr %5.1f%% samples  %s" % (100 * c / tot, 100 * bw[f] / totw, f))
for f, r in sorted(lines, key=lambda fr: -int(fr[1][iW] or 0))[:top_n]:
    print("%5.2f%% smp %5.2f%% ins  %s:%s  %s" % (100 * int(r[iW] or 0) / totw, 100 * int(r[iI] or 0) / tot, f, r[0],
                                                 r[1].strip()[:120]))
