#!/usr/bin/env python3
"""BASELINE config 5: every reference scene that loads, 7680x4320, 1..N GPUs.

  python tools/config5.py [--frames K] [--width W --height H] [--mode scenes|tiles]
  python -m torch.distributed.run --nproc-per-node N ... tools/config5.py --mode tiles

mode scenes: whole scenes are dealt round-robin to the ranks (replicas, no collective at all);
mode tiles : every rank renders its round-robin 16-row strips of every scene.
Prints one JSON line: per-scene ms/frame and Mrays/s, aggregate Mrays/s (segments of all scenes /
max over ranks of the device time), and -- on 1 GPU -- the reference kernel (G0) on the same frames.
Scenes come from tests/golden (crate_land only when its fixture was generated locally:
tests/golden/make_fixtures.py).  Cubemaps: the reference's shipped 4096x3072 crosses for the scenes
that name them (color_sample / crate_land: field_with_house.jpg, island: water.jpg -> six 1024^2
faces, 100 MB) when tests/golden/_large was generated (tests/golden/make_cubemap_fixture.py); a
smooth synthetic cubemap otherwise and for the scenes whose JPEG is not shipped (same for both
implementations; the line says which)."""
import argparse, copy, json, os, sys
import numpy as np
import torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "tests"))
import bench
from pt_testlib import load_scene, load_cube_cross, RefGpu

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=8)
ap.add_argument("--width", type=int, default=7680)
ap.add_argument("--height", type=int, default=4320)
ap.add_argument("--bounces", type=int, default=3)
ap.add_argument("--mode", default="tiles", choices=["scenes", "tiles"])
ap.add_argument("--no-reference", action="store_true")
ap.add_argument("--out", default=None, help="also write the JSON line to this file (stdout carries NCCL's banner too)")
a = ap.parse_args()

rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
torch.cuda.set_device(local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
P = bench.pkg()
names = ["indoor_tex", "crate_land_plain", "island_plain", "color_sample_plain", "sss_crate_plain"]
scenes, cubes = {}, {}
for n in names:
    try:
        sc = copy.copy(load_scene(n))
    except Exception:
        continue
    cross = {"color_sample_plain": "field_with_house", "crate_land_plain": "field_with_house", "island_plain": "water"}.get(n)
    img = load_cube_cross(cross) if cross else None
    if img is not None:
        faces, _ = P.core.cubemap_faces_from_cross(img)
        sc.cubemap, sc.cubemap_size = faces, faces.shape[1]
        cubes[n] = f"{cross}.jpg ({faces.shape[1]}^2 faces)"
    elif sc.cubemap is None or n != "indoor_tex":
        sc.cubemap, sc.cubemap_size = P.synth.gradient_cubemap(32), 32
        cubes[n] = "synthetic gradient 32^2"
    else:
        cubes[n] = "constant colour (garden.jpg is not shipped: the reference's fallback)"
    if n.startswith("sss_crate"):  # legacy camera line -> fov 0 in the reference's parser
        sc.camera = P.synth.camera_array((-2.7, 2.06, 2.52), (0.62, -0.348, -0.7), aperture=0.02, focus_dist=3.555)
    scenes[n] = sc
core = P.Core(local)
w, h = a.width, a.height
gather = P.tiles.BandGather(h, w, world, rank, interleaved=True)
per_scene, total_ms, total_seg = {}, 0.0, 0
dev = torch.device("cuda", local)
for i, (n, sc) in enumerate(scenes.items()):
    mine = (i % world == rank) if a.mode == "scenes" else True
    rows_kw = {"strips": gather.strips} if (a.mode == "tiles" and world > 1) else {}
    bh = gather.n_rows if rows_kw else h
    ms, seg = 0.0, 0
    if mine:
        sid = core.upload_scene(sc); cid = core.set_cubemap(sc.cubemap)
        acc = torch.zeros(bh, w, 3, device=dev); out = torch.zeros(bh, w, dtype=torch.int32, device=dev)
        for k in (1, 2):
            core.render(sid, cid, sc, w, h, acc, out, None, bounces=a.bounces, frame_nb=k, **rows_kw)
        torch.cuda.synchronize(); core.reset_counters()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(3, 3 + a.frames):
            core.render(sid, cid, sc, w, h, acc, out, None, bounces=a.bounces, frame_nb=k, **rows_kw)
        e1.record(); torch.cuda.synchronize()
        ms, seg = e0.elapsed_time(e1), core.counters().segments
        del acc, out
    per_scene[n] = (ms, seg)
    total_ms += ms; total_seg += seg
t = torch.tensor([total_ms, float(total_seg)] + [v for n in scenes for v in per_scene[n]], dtype=torch.float64, device=dev)
mx = t.clone()
if world > 1:
    dist.all_reduce(mx, op=dist.ReduceOp.MAX); dist.all_reduce(t, op=dist.ReduceOp.SUM)
if rank == 0:
    line = {"config": f"BASELINE configs[4] (config 5): {len(scenes)} scenes {w}x{h} bounces={a.bounces} {a.frames} frames each",
            "mode": a.mode if world > 1 else "one GPU renders every scene", "n_gpus": world,
            "aggregate_Mrays_s": float(t[1]) / (float(mx[0]) * 1e-3) / 1e6, "ms_all_scenes_max_rank": float(mx[0]), "scenes": {}}
    for j, n in enumerate(scenes):
        ms, seg = float(mx[2 + 2 * j]), float(t[3 + 2 * j])
        line["scenes"][n] = {"faces": scenes[n].n_faces, "cubemap": cubes.get(n), "ms_per_frame": ms / a.frames,
                             "Mrays_s": seg / (ms * 1e-3) / 1e6 if ms else None}
    if world == 1 and not a.no_reference:
        ref = {}
        for n, sc in scenes.items():
            try:
                r = bench.reference_kernel_on_gpu(sc, w, h, a.bounces, 2, 1)
                ref[n] = {"ms_per_frame": r["ms_per_step"], "Mrays_s": r["value"],
                          "speedup": r["ms_per_step"] / line["scenes"][n]["ms_per_frame"]}
            except Exception as e:
                ref[n] = {"error": str(e)}
        line["reference_kernel_same_gpu"] = ref
    print(json.dumps(line))
    if a.out:
        with open(a.out, "w") as f:
            f.write(json.dumps(line, indent=1) + "\n")
core.close()
if world > 1:
    dist.destroy_process_group()
