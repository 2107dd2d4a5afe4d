#!/usr/bin/env python3
"""Iterations in the first stage / in later stages, by frame size and bounce count (ms/frame as rendered).
usage: sweep_first.py <scene> [split_extend=-1]"""
import json, os, sys
import torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import bench
P = bench.pkg()
name = sys.argv[1]
form = int(sys.argv[2]) if len(sys.argv) > 2 else -1
sc = bench.load_scene(P, name)
core = P.Core(0); sid = core.upload_scene(sc); cid = core.set_cubemap(sc.cubemap)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
sizes = [(640, 360), (1280, 720), (1920, 1080), (2720, 1530), (3840, 2160)] if not name.startswith("synth") else [(1920, 1080)]
for (w, h) in sizes:
    acc = torch.zeros(h, w, 3, device="cuda"); out = torch.zeros(h, w, dtype=torch.int32, device="cuda")
    for b in ((3, 4, 8) if not name.startswith("synth") else (4,)):
        res = {"scene": name, "w": w, "h": h, "b": b}
        n = max(8, min(64, int(40e6 / (w * h))))
        for tag, (first, nxt) in {"1/1": (1, 1), "2/1": (2, 1), "2/2": (2, 2), "3/1": (3, 1)}.items():
            core.set_wavefront_auto()
            core.set_wavefront(first, nxt, form, 0)
            for k in range(1, 4):
                core.render(sid, cid, sc, w, h, acc, out, None, bounces=b, frame_nb=k)
            torch.cuda.synchronize()
            best = 1e9
            for rep in range(3):
                e0.record()
                for k in range(4, 4 + n):
                    core.render(sid, cid, sc, w, h, acc, out, None, bounces=b, frame_nb=k)
                e1.record(); torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1) / n)
            res[tag] = round(best, 4)
        print(json.dumps(res), flush=True)
