#!/usr/bin/env python3
"""Fused vs split later stages as a function of scene size (development aid)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
P = bench.pkg()
w, h, b = 1920, 1080, 4
core = P.Core(0)
acc = torch.zeros(h, w, 3, device="cuda"); out = torch.zeros(h, w, dtype=torch.int32, device="cuda")
for grid in (24, 50, 100, 150, 200, 300, 500):
    sc = P.synth.terrain_room(grid)
    sid = core.upload_scene(sc); cid = core.set_cubemap(sc.cubemap)
    info = core.scene_info(sid)
    kb = (info.bvh_nodes * info.bvh_node_bytes + info.n_faces * info.tri_bytes) / 1024
    res = []
    for split in (0, 1):
        core.set_wavefront(1, 1, split, 12)
        for k in range(1, 4):
            core.render(sid, cid, sc, w, h, acc, out, None, bounces=b, frame_nb=k)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(4, 24):
            core.render(sid, cid, sc, w, h, acc, out, None, bounces=b, frame_nb=k)
        e1.record(); torch.cuda.synchronize()
        res.append(e0.elapsed_time(e1) / 20)
    print(f"grid {grid}: {info.n_faces} faces, {kb:.0f} KB: fused {res[0]:.4f} split {res[1]:.4f} ms/frame")
