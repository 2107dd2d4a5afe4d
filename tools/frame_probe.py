#!/usr/bin/env python3
"""Renders a few frames of the headline workload (for ncu launch lists / captures)."""
import os, sys
import torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import bench
P = bench.pkg()
name = sys.argv[1] if len(sys.argv) > 1 else "indoor_tex"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 6
w, h, b = (int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])) if len(sys.argv) > 5 else (1920, 1080, 3)
sc = bench.load_scene(P, name)
core = P.Core(0); sid = core.upload_scene(sc); cid = core.set_cubemap(sc.cubemap)
acc = torch.zeros(h, w, 3, device="cuda"); out = torch.zeros(h, w, dtype=torch.int32, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for k in range(1, 4):
    core.render(sid, cid, sc, w, h, acc, out, None, bounces=b, frame_nb=k)
torch.cuda.synchronize()
e0.record()
for k in range(4, 4 + frames):
    core.render(sid, cid, sc, w, h, acc, out, None, bounces=b, frame_nb=k)
e1.record(); torch.cuda.synchronize()
print(f"{name} {w}x{h} b={b}: {e0.elapsed_time(e1)/frames:.4f} ms/frame")
