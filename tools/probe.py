#!/usr/bin/env python3
"""A/B probe: ms/frame of a workload as rendered (two half-frame pipelines) + per-launch event
times and traversal counters from ptc_profile_frame (one pipeline).
usage: probe.py <scene> <frames> <w> <h> <bounces>   (PTC_LIBRARY / PTC_* env select the variant)"""
import os, sys, json, time
import torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import bench
P = bench.pkg()
name = sys.argv[1] if len(sys.argv) > 1 else "indoor_tex"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 16
w, h, b = (int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])) if len(sys.argv) > 5 else (1920, 1080, 3)
t0 = time.time()
sc = bench.load_scene(P, name)
t1 = time.time()
core = P.Core(0); sid = core.upload_scene(sc); cid = core.set_cubemap(sc.cubemap)
t2 = time.time()
acc = torch.zeros(h, w, 3, device="cuda"); out = torch.zeros(h, w, dtype=torch.int32, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for k in range(1, 4):
    core.render(sid, cid, sc, w, h, acc, out, None, bounces=b, frame_nb=k)
torch.cuda.synchronize()
best = 1e9
for rep in range(3):
    e0.record()
    for k in range(4, 4 + frames):
        core.render(sid, cid, sc, w, h, acc, out, None, bounces=b, frame_nb=k)
    e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1) / frames)
pr = core.profile_frame(sid, cid, sc, w, h, acc, out, None, bounces=b, frame_nb=5, repeats=4)
kinds = {}
for k, ms in zip(pr["launch_kind"], pr["launch_ms"]):
    kinds[k] = kinds.get(k, 0.0) + ms
tag = os.environ.get("PTC_TAG", os.path.basename(os.environ.get("PTC_LIBRARY", "default")))
print(json.dumps({"tag": tag, "scene": name, "w": w, "h": h, "b": b, "ms_frame": round(best, 4),
                  "ms_by_kind": {k: round(v, 4) for k, v in kinds.items()},
                  "nodes_per_ray": round(pr["nodes_visited"] / max(1, pr["traced"]), 3),
                  "tris_per_ray": round(pr["tris_tested"] / max(1, pr["traced"]), 3),
                  "traced": pr["traced"], "ms_frame_1pipe": round(pr["frame_ms"], 4),
                  "extend_lanes_per_iter": [round(x / max(1, pr["extend_lanes"][0]), 2) for x in pr["extend_lanes"][1:]],
                  "extend_iters": pr["extend_lanes"][0], "load_s": round(t1 - t0, 1), "upload_s": round(t2 - t1, 1),
                  "checksum": int(acc.view(torch.int32).sum(dtype=torch.int64).item())}))
