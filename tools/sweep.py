#!/usr/bin/env python3
"""Sweeps the wavefront schedule knobs on the headline workload (development aid)."""
import os, sys
import torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import bench
P = bench.pkg()
sc = bench.load_scene(P, sys.argv[1] if len(sys.argv) > 1 else "indoor_tex")
core = P.Core(0); sid = core.upload_scene(sc); cid = core.set_cubemap(sc.cubemap)
w, h, b = 1920, 1080, int(sys.argv[2]) if len(sys.argv) > 2 else 3
acc = torch.zeros(h, w, 3, device="cuda"); out = torch.zeros(h, w, dtype=torch.int32, device="cuda")
def run(label):
    for k in range(1, 4):
        core.render(sid, cid, sc, w, h, acc, out, None, bounces=b, frame_nb=k)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(4, 24):
        core.render(sid, cid, sc, w, h, acc, out, None, bounces=b, frame_nb=k)
    e1.record(); torch.cuda.synchronize()
    pr = core.profile_frame(sid, cid, sc, w, h, acc, out, None, bounces=b, frame_nb=5, repeats=5)
    print(f"{label:32s} {e0.elapsed_time(e1)/20:.4f} ms/frame  launches " +
          " ".join(f"{k[0]}{ms*1e3:.0f}" for k, ms in zip(pr['launch_kind'], pr['launch_ms'])))
for first, nxt, split, refill in [(1,1,0,8),(2,1,0,8),(3,1,0,8),(1,1,1,1),(1,1,1,4),(1,1,1,8),(1,1,1,12),(1,1,1,16),(1,1,1,24),(1,1,1,32),(2,1,1,8)]:
    core.set_wavefront(first, nxt, split, refill)
    run(f"first={first} next={nxt} split={split} refill={refill}")
