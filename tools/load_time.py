#!/usr/bin/env python3
"""Load wall-clock of the 10 M-triangle scene: indexed mesh -> Face[] (ptc_flatten_faces,
the parallel form of scene.cpp:202-283), BVH build (ptc_host_bvh_build) and the whole
ptc_upload_scene_indexed (flatten + build + record split + host->device copies).
usage: load_time.py [grid=2236] [out.json]"""
import json, os, sys, time
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import bench
P = bench.pkg()
grid = int(sys.argv[1]) if len(sys.argv) > 1 else 2236
t0 = time.time()
sc = P.synth.terrain_room(grid=grid, tex_size=256)
t_gen = time.time() - t0
f = np.frombuffer(sc.faces, dtype=np.float32).reshape(-1, 28) if not isinstance(sc.faces, np.ndarray) else sc.faces.view(np.float32).reshape(-1, 28)
n = f.shape[0]
ar = np.arange(n * 3, dtype=np.int32)
idx = np.stack([ar, ar, ar], axis=1)
mesh = P.ptscene.IndexedMesh(f[:, 0:9].reshape(-1, 3), f[:, 9:18].reshape(-1, 3), f[:, 18:24].reshape(-1, 2),
                             [(idx, f[:, 27].view(np.int32).copy())], stride=3)
res = {"faces": int(n), "host_threads": os.cpu_count(), "generate_s": round(t_gen, 2)}
for threads in (None, "1"):
    if threads:
        os.environ["PTC_BUILD_THREADS"] = threads
    t0 = time.time(); faces, counts = P.core.flatten_faces(mesh); dt = time.time() - t0
    res["flatten_s" + ("_1thread" if threads else "")] = round(dt, 3)
    os.environ.pop("PTC_BUILD_THREADS", None)
same = bool((faces.view(np.uint32).reshape(-1, 28)[:, :24] == f.view(np.uint32)[:, :24]).all())
res["flatten_equals_source_faces"] = same
t0 = time.time(); b = P.host_bvh(sc); res["bvh_build_s"] = round(time.time() - t0, 2)
res["bvh_nodes"] = int(len(b["nodes"])) if isinstance(b, dict) and "nodes" in b else None
import torch
if torch.cuda.is_available():
    core = P.Core(0)
    t0 = time.time(); sid = core.upload_scene_indexed(mesh, sc); torch.cuda.synchronize()
    res["upload_scene_indexed_s"] = round(time.time() - t0, 2)
    t0 = time.time(); sid2 = core.upload_scene(sc); torch.cuda.synchronize()
    res["upload_scene_s"] = round(time.time() - t0, 2)
    core.close()
print(json.dumps(res))
if len(sys.argv) > 2:
    json.dump(res, open(sys.argv[2], "w"), indent=1)
