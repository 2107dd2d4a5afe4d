#!/usr/bin/env python3
"""Renders exactly F frames of a workload through ptc_render (nothing else launches stage kernels):
the program ncu counts per-frame instructions / DRAM bytes on (tools/ncu_counters.py).
usage: frames_only.py <scene> <frames> <w> <h> <bounces>"""
import os, sys
import torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import bench
P = bench.pkg()
name, frames, w, h, b = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
sc = bench.load_scene(P, name)
core = P.Core(0); sid = core.upload_scene(sc); cid = core.set_cubemap(sc.cubemap)
acc = torch.zeros(h, w, 3, device="cuda"); out = torch.zeros(h, w, dtype=torch.int32, device="cuda")
for k in range(1, frames + 1):
    core.render(sid, cid, sc, w, h, acc, out, None, bounces=b, frame_nb=k)
torch.cuda.synchronize()
print(f"{name} {w}x{h} b={b}: {frames} frames, {core.launch_count()} launches")
