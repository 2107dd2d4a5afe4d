#!/usr/bin/env python3
"""Basic-block level view of an .ncu-rep's SASS page: instructions executed, lanes, samples."""
import csv, subprocess, sys
rep = sys.argv[1]; which = int(sys.argv[2]) if len(sys.argv) > 2 else 0; thresh = float(sys.argv[3]) if len(sys.argv) > 3 else 0.006
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--print-kernel-base", "function"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
ks = []; cur = None
for r in rows:
    if r and r[0] == "Kernel Name": cur = []; ks.append(cur); continue
    if r and r[0] == "Address": hdr = r; continue
    if cur is not None and len(r) > 10: cur.append(r)
ci = hdr.index("Instructions Executed"); ti = hdr.index("Thread Instructions Executed"); si = hdr.index("# Samples")
k = ks[which]
tot_i = sum(int(r[ci]) for r in k); tot_t = sum(int(r[ti]) for r in k); tot_s = sum(int(r[si]) for r in k)
print(f"kernel {which}: rows {len(k)} warp-inst {tot_i} thread-inst {tot_t} lanes/inst {tot_t/tot_i:.1f} samples {tot_s}")
start = 0
for i in range(1, len(k) + 1):
    if i == len(k) or k[i][ci] != k[start][ci] or 'BRA' in k[i-1][1] or 'BSYNC' in k[i-1][1]:
        n = i - start; ie = int(k[start][ci]); te = sum(int(r[ti]) for r in k[start:i]); sm = sum(int(r[si]) for r in k[start:i])
        if ie * n > tot_i * thresh or sm > tot_s * thresh:
            ops = {}
            for r in k[start:i]:
                t = r[1].split(); op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]; ops[op] = ops.get(op, 0) + 1
            top = sorted(ops.items(), key=lambda x: -x[1])[:6]
            print(f"@{start:5d} n={n:4d} exec={ie:8d} warpinst={ie*n/tot_i*100:5.1f}% lanes={te/max(1,ie*n):5.1f} samples={sm/tot_s*100:5.1f}% {top}")
        start = i
