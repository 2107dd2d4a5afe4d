#!/usr/bin/env python3
"""cuobjdump summary of libptcore.so: per kernel registers / stack (spill) / shared memory, SASS
instruction count, counts of the sm_100-specific instructions DESIGN.md 4 names, spill instructions,
and one sample line of each.  usage: sass_excerpt.py [lib] > profiles/rN_sass_excerpt.txt"""
import os, re, subprocess, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(REPO, "cuda-pathtracer_b200", "libptcore.so")
res = subprocess.run(["cuobjdump", "-res-usage", lib], capture_output=True, text=True).stdout
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
usage = {}
cur = None
for line in res.splitlines():
    m = re.search(r"Function (\S+):", line)
    if m:
        cur = m.group(1)
    elif cur and "REG:" in line:
        usage[cur] = line.strip()
        cur = None
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip().split("(")[0]
WATCH = ["LDG.E.ENL2.256", "FFMA2", "HADD2.F32", "LDG.E.128", "LDL", "STL", "LDS", "STS", "CCTL", "REDUX", "MATCH", "VOTE", "SHFL"]
print(f"# {os.path.relpath(lib, REPO)}: cuobjdump -res-usage / -sass summary (sm_100a)\n")
blocks = re.split(r"\n\s*Function : ", sass)[1:]
for blk in blocks:
    name = blk.split("\n", 1)[0].strip()
    if not any(k in name for k in ("k_path", "k_extend", "k_wave", "k_primary", "k_intersect")):
        continue
    ins = re.findall(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_.]+)[ ;]", blk, flags=re.M)
    print(f"## {demangle(name)}\n   {usage.get(name, '')}\n   SASS instructions: {len(ins)}")
    counts = {w: sum(1 for i in ins if i.startswith(w)) for w in WATCH}
    print("   " + "  ".join(f"{w}: {c}" for w, c in counts.items() if c))
    for w in ("LDG.E.ENL2.256", "FFMA2", "HADD2.F32"):
        m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(.*?%s.*?);" % re.escape(w), blk, flags=re.M)
        if m:
            print(f"   e.g. {m.group(1).strip()}")
    print()
