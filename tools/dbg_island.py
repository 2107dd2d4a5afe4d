import os, sys
import numpy as np, torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "tests"))
from pt_testlib import Oracle, RefGpu, load_scene, pkg, cube_of
P = pkg()
name = sys.argv[1] if len(sys.argv) > 1 else "island_plain"
sc = load_scene(name)
if sc.cubemap is None:
    sc.cubemap = P.synth.gradient_cubemap(8); sc.cubemap_size = 8
core = P.Core(0); sid = core.upload_scene(sc); cid = core.set_cubemap(cube_of(sc)[0])
g0 = RefGpu(sc); orc = Oracle()
w, h = 320, 180
ids, t = core.trace_primary(sid, sc, w, h, 1)
mats = sc.material_ids()
for b in (1, 2, 3):
    acc_m = torch.zeros(h, w, 3, device="cuda"); out_m = torch.zeros(h, w, dtype=torch.int32, device="cuda"); rad_m = torch.zeros(h, w, 3, device="cuda")
    acc_g = torch.zeros(h, w, 3, device="cuda"); out_g = torch.zeros(h, w, dtype=torch.int32, device="cuda")
    core.render(sid, cid, sc, w, h, acc_m, out_m, rad_m, bounces=b, frame_nb=1)
    g0.kernel(w, h, 1, b, acc_g.data_ptr(), out_g.data_ptr())
    torch.cuda.synchronize()
    am, ag = acc_m.cpu().numpy()[::-1], acc_g.cpu().numpy()[::-1]
    neq = (am.view(np.uint32) != ag.view(np.uint32)).any(-1)
    print("bounces", b, "differ", int(neq.sum()), "max abs", float(np.nanmax(np.abs(am - ag))), "mean abs", float(np.nanmean(np.abs(am-ag))))
    ref = orc.radiance(sc, w, h, 1, b)
    rm = rad_m.cpu().numpy()
    print("   vs oracle unclamped: max abs", float(np.nanmax(np.abs(rm-ref))), "frac close 1e-3", float(np.isclose(rm, ref, rtol=1e-3, atol=1e-4).all(-1).mean()))
    for mid in range(len(sc.materials)):
        sel = (ids >= 0) & (ids < sc.n_faces)
        m = np.zeros_like(sel); m[sel] = mats[ids[sel]] == mid
        print("   material", mid, sc.materials[mid], "pixels", int(m.sum()), "differ", int(neq[m].sum()))
    m = ids < 0
    print("   miss pixels", int(m.sum()), "differ", int(neq[m].sum()))
    ys, xs = np.nonzero(neq)
    for k in range(min(3, len(ys))):
        y, x = ys[k], xs[k]
        print("   px", x, y, "id", ids[y, x], "mine", am[y, x], "g0", ag[y, x], "oracle", ref[y, x])
