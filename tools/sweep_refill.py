import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
P = bench.pkg()
sc = bench.load_scene(P, sys.argv[1])
b, w, h = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
core = P.Core(0); sid = core.upload_scene(sc); cid = core.set_cubemap(sc.cubemap)
acc = torch.zeros(h, w, 3, device="cuda"); out = torch.zeros(h, w, dtype=torch.int32, device="cuda")
for refill in (6, 8, 10, 12, 14, 16, 20):
    core.set_wavefront(1, 1, 1, refill)
    for k in range(1, 4):
        core.render(sid, cid, sc, w, h, acc, out, None, bounces=b, frame_nb=k)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(4, 24):
        core.render(sid, cid, sc, w, h, acc, out, None, bounces=b, frame_nb=k)
    e1.record(); torch.cuda.synchronize()
    print(f"{sys.argv[1]} refill={refill}: {e0.elapsed_time(e1)/20:.4f} ms/frame")
