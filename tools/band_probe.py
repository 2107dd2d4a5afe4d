#!/usr/bin/env python3
"""What one rank of a strong-scaled frame costs: strips (16, N, 0) of a W x H frame on ONE GPU, under
different wavefront granularities (iterations per later stage).  usage: band_probe.py W H bounces N"""
import json, os, sys
import torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import bench
P = bench.pkg()
w, h, b, n = (int(x) for x in sys.argv[1:5])
sc = bench.load_scene(P, "indoor_tex")
core = P.Core(0); sid = core.upload_scene(sc); cid = core.set_cubemap(sc.cubemap)
strips = (16, n, 0)
bh = core.band_rows(w, h, strips=strips)
acc = torch.zeros(bh, w, 3, device="cuda"); out = torch.zeros(bh, w, dtype=torch.int32, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
res = {"w": w, "h": h, "bounces": b, "ranks": n, "band_rows": bh}
for tag, (first, nxt) in {"1/1": (1, 1), "1/2": (1, 2), "1/3": (1, 3), "1/4": (1, 4), "2/2": (2, 2), "2/3": (2, 3), "1/8": (1, 8)}.items():
    core.set_wavefront(first, nxt, 0, 0)
    for k in range(1, 4):
        core.render(sid, cid, sc, w, h, acc, out, None, bounces=b, frame_nb=k, strips=strips)
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(3):
        e0.record()
        for k in range(4, 36):
            core.render(sid, cid, sc, w, h, acc, out, None, bounces=b, frame_nb=k, strips=strips)
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / 32)
    res[tag] = round(best, 4)
core.set_wavefront_auto()
print(json.dumps(res))
