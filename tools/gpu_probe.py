#!/usr/bin/env python3
"""Development probe (GPU box): product vs G0 (reference kernel) vs oracle, plus timings."""
import json
import os
import sys
import time

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "tests"))
from pt_testlib import Oracle, RefGpu, load_scene, pkg, cube_of, GOLDEN  # noqa: E402

P = pkg()
name = sys.argv[1] if len(sys.argv) > 1 else "indoor_tex"
sc = load_scene(name)
core = P.Core(0)
sid = core.upload_scene(sc)
cid = core.set_cubemap(cube_of(sc)[0])
g0 = RefGpu(sc)
orc = Oracle()
info = core.scene_info(sid)
print("scene", name, "faces", info.n_faces, "nodes", info.bvh_nodes, "tex bytes", info.texture_bytes)

# 1. RNG
kat = json.load(open(os.path.join(GOLDEN, "rng_kat.json")))
seeds = np.array([int(s) for s in kat["uniforms"]], np.uint32)
want = np.array([[float.fromhex(v) for v in kat["uniforms"][s]] for s in kat["uniforms"]], np.float32)
mine = core.debug_rng(seeds, 8)
cur = core.debug_rng(seeds, 8, reference_impl=True)
print("rng: mine==KAT", (mine == want).all(), " curand(device)==KAT", (cur == want).all())

# 2. primary hits
w, h = 640, 360
ids, t = core.trace_primary(sid, sc, w, h, 1)
o_g, d_g, ids_g, t_g = g0.primary(w, h, 1)
o_c, d_c = orc.primary_rays(sc, w, h, 1)
ids_c, t_c = orc.intersect(sc, o_c, d_c)
print("primary 640x360: vs G0 id mismatches", int((ids != ids_g).sum()), "t bit mismatches",
      int((t.view(np.uint32) != t_g.view(np.uint32)).sum()),
      "| vs oracle id mismatches", int((ids.reshape(-1) != ids_c).sum()),
      "| G0 vs oracle", int((ids_g.reshape(-1) != ids_c).sum()))
print("  ray bits G0 vs oracle: o", int((o_g.view(np.uint32) != o_c.view(np.uint32)).sum()),
      "d", int((d_g.view(np.uint32) != d_c.view(np.uint32)).sum()), "of", o_g.size)

# 3. random rays
rng = np.random.default_rng(1)
v = sc.faces_f32()[:, :9].reshape(-1, 3)
lo, hi = v.min(0), v.max(0)
n = 2_000_000
o = (lo + (hi - lo) * rng.random((n, 3))).astype(np.float32)
d = rng.normal(size=(n, 3)).astype(np.float32)
d /= np.linalg.norm(d, axis=1, keepdims=True)
core.reset_counters()
i1, t1 = core.intersect_rays(sid, o, d)
c = core.counters()
i2, t2 = g0.intersect(o, d)
print("random rays 2M: id mismatches", int((i1 != i2).sum()), "t bit mismatches",
      int((t1.view(np.uint32) != t2.view(np.uint32)).sum()), "hit frac", float((i2 >= 0).mean()),
      "nodes/ray", c.nodes_visited / n, "tris/ray", c.tris_tested / n)

# 4. radiance vs G0, one frame
for (w, h, b) in [(160, 90, 3), (640, 360, 4)]:
    acc_m = torch.zeros(h, w, 3, device="cuda")
    out_m = torch.zeros(h, w, dtype=torch.int32, device="cuda")
    rad_m = torch.zeros(h, w, 3, device="cuda")
    core.render(sid, cid, sc, w, h, acc_m, out_m, rad_m, bounces=b, frame_nb=1)
    acc_g = torch.zeros(h, w, 3, device="cuda")
    out_g = torch.zeros(h, w, dtype=torch.int32, device="cuda")
    g0.kernel(w, h, 1, b, acc_g.data_ptr(), out_g.data_ptr())
    torch.cuda.synchronize()
    am, ag = acc_m.cpu().numpy(), acc_g.cpu().numpy()
    neq = (am.view(np.uint32) != ag.view(np.uint32)).any(axis=2)
    print(f"accum {w}x{h} b={b}: pixels differing from G0 (bitwise) {int(neq.sum())}/{w*h}; "
          f"max abs diff {float(np.abs(am-ag).max()):.3e}; mean abs {float(np.abs(am-ag).mean()):.3e}; "
          f"rgba differing {int((out_m != out_g).sum())}")
    if (w, h, b) == (160, 90, 3):
        z = np.load(os.path.join(GOLDEN, f"ref_{name}_160x90.npz")) if os.path.exists(os.path.join(GOLDEN, f"ref_{name}_160x90.npz")) else None
        if z is not None:
            ref = z["radiance_b3_f1"]
            rm = rad_m.cpu().numpy()
            close = np.isclose(rm, ref, rtol=1e-3, atol=1e-4).all(axis=2)
            print(f"  vs golden (reference on host cores): pixels within 1e-3: {int(close.sum())}/{w*h}; "
                  f"mean image diff {float(np.abs(rm.mean((0,1))-ref.mean((0,1))).max()):.3e}")

# 5. timing
w, h, b = 1920, 1080, 3
acc = torch.zeros(h, w, 3, device="cuda")
out = torch.zeros(h, w, dtype=torch.int32, device="cuda")
for label, fn in [("product", lambda k: core.render(sid, cid, sc, w, h, acc, out, None, bounces=b, frame_nb=k)),
                  ("G0", lambda k: g0.kernel(w, h, k, b, acc.data_ptr(), out.data_ptr()))]:
    for k in range(1, 4):
        fn(k)
    torch.cuda.synchronize()
    core.reset_counters()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    nfr = 20
    e0.record()
    for k in range(1, nfr + 1):
        fn(k)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / nfr
    c = core.counters()
    segs = c.segments / nfr if label == "product" else None
    print(f"timing {label}: {ms:.3f} ms/frame @1080p b={b}" + (f"; segments/frame {segs:.0f} -> {segs/ms/1e3:.1f} Mrays/s; per bounce {[c.rays_per_bounce[i]//nfr for i in range(b)]}" if segs else ""))
