#!/usr/bin/env python3
"""bench.py -- BASELINE.json's metric on BASELINE.json's config.

    python bench.py --gpus N --steps K --warmup W            # this repository's CUDA path
    python bench.py --impl reference --gpus N --steps K ...   # the reference on host cores

metric    Mrays/s @1080p, 1 spp/frame, README indoor scene (frames/s reported alongside).
          A ray = one radiance-loop iteration = one reference intersect() call (SURVEY 8d
          "segments"); identical for both implementations under identical seeds.
workload  BASELINE.json configs[1]: indoor scene (textures as the MTL names them, cubemap =
          the reference's constant fallback since garden.jpg is not shipped), 1920x1080,
          1 spp per frame, temporal accumulation, bounces = 3 (the reference's constant).
step      one frame = one raytrace() call = 1 spp over the whole image.
N > 1     weak scaling by image tiles: the image grows to N x 1080p pixels of the same 16:9 view
          (2720x1530, 3840x2160, 5440x3060), the scene is replicated, every rank accumulates its
          own round-robin 16-row strips, and the full RGBA8 image is assembled on rank 0 EVERY frame
          by the kernels' own stores over NVLink (--gather store, default; no collective), inside the
          per-frame timing.  --gather nccl: one all_gather_into_tensor after the last timed frame.
value     device-resident: accumulator and RGBA8 stay in HBM, CUDA events per frame on the
          launching stream, L2 flushed (512 MB memset) between timed frames.
e2e       the same frames through ptc_render_host: camera/arguments host->device and the
          RGBA8 frame device->pinned host every step, inside the timed region.
roofline  for the frame's stage kernels: three lower bounds on their time -- bytes (BVH bytes at
          the MEASURED bandwidth of the cache tier the BVH lives in + every other byte at the HBM
          peak of MEASURED_PEAKS.json), intersection flops at the FP32 peak, and warp instructions
          at the issue rate (4 schedulers x SMs x the SM clock sampled during the run) -- `bound`
          names the largest, `frac` = that bound / the kernels' CUDA-event time.  Instruction
          counts and DRAM traffic come from the committed ncu capture of the same workload
          (profiles/r2_counters.json), and say so.
sustained >= 5 s of back-to-back frames (no flush) with the SM clock sampled during them.
workloads --workload headline (default; BASELINE.json configs[1]) | cfg0 | cfg1b4 | cfg2 | cfg3:
          configs[0] 640x360 b=4, configs[1] at 4 iterations, configs[2] = 3840x2160, 16 frames,
          8 iterations, STRONG scaling over the ranks with the gathered accumulator compared bit
          for bit with a 1-GPU render, configs[3] = the 10 M-triangle synthetic mesh.
"""
from __future__ import annotations

import argparse
import ctypes as C
import importlib.util
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

REPO = os.path.dirname(os.path.abspath(__file__))
PKG_DIR = os.path.join(REPO, "cuda-pathtracer_b200")
GOLDEN = os.path.join(REPO, "tests", "golden")
# weak scaling: N x the pixels of the SAME 16:9 view (the camera frames the same picture, so the
# cost per ray does not change with N); 4 GPUs = BASELINE config 3's 4K frame
WEAK_DIMS = {1: (1920, 1080), 2: (2720, 1530), 4: (3840, 2160), 8: (5440, 3060)}


def pkg():
    name = "cuda_pathtracer_b200"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, os.path.join(PKG_DIR, "__init__.py"),
                                                  submodule_search_locations=[PKG_DIR])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def load_scene(P, name):
    if name.startswith("synth"):
        return P.synth.terrain_room(grid=int(name[5:] or 256), tex_size=256)
    p = os.path.join(GOLDEN, name + ".ptscene")
    if not os.path.exists(p):
        p += ".xz"
    sc = P.ptscene.load_ptscene(p)
    if sc.cubemap is None:
        sc.cubemap = P.synth.constant_cubemap((19 / 255.0, 27 / 255.0, 35 / 255.0))
        sc.cubemap_size = 1
    return sc


def measured_peaks():
    p = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = [r for (t, r) in self.rows if t0 - 0.05 <= t <= t1 + 0.15] or [r for (_, r) in self.rows]
        for r in rows:
            f = [x.strip() for x in r.split(",")]
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except Exception:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------
# the reference arm / cpu_baseline: the reference's own code (oracle/_ref) or the oracle port
# ---------------------------------------------------------------------------------------------
class CpuRenderer:
    def __init__(self, prefer_reference=True):
        ref = os.path.join(REPO, "oracle", "_ref", "libref_cpu.so")
        port = os.path.join(REPO, "oracle", "libpt_oracle.so")
        self.orc = C.CDLL(port) if os.path.exists(port) else None
        if prefer_reference and os.path.exists(ref):
            self.lib, self.kind, self.fn = C.CDLL(ref), "reference", "ref_cpu_kernel_sample"
        elif self.orc is not None:
            self.lib, self.kind, self.fn = self.orc, "port", "oracle_kernel_sample"
        else:
            raise RuntimeError("neither oracle/_ref/libref_cpu.so nor oracle/libpt_oracle.so is built")
        self.cores = os.cpu_count() or 1

    def count(self, sc, w, h, frame, bounces, stride, offset):
        """Pixels and segments of the sample (untimed; the oracle port counts them)."""
        desc, cam = sc.to_desc(), sc.camera_struct()
        acc = np.zeros((h, w, 3), np.float32)
        px, seg = C.c_uint64(0), C.c_uint64(0)
        self.orc.oracle_kernel_sample(C.byref(desc), sc.cubemap.ctypes.data_as(C.c_void_p), C.c_uint(sc.cubemap_size),
                                      C.byref(cam), C.c_uint(w), C.c_uint(h), C.c_uint(frame), C.c_uint(bounces),
                                      C.c_uint(stride), C.c_uint(offset), acc.ctypes.data_as(C.c_void_p), None,
                                      C.byref(px), C.byref(seg), C.c_int(self.cores))
        return px.value, seg.value

    def run(self, sc, w, h, frame, bounces, stride, offset, acc, out):
        desc, cam = sc.to_desc(), sc.camera_struct()
        args = [C.byref(desc), sc.cubemap.ctypes.data_as(C.c_void_p), C.c_uint(sc.cubemap_size), C.byref(cam),
                C.c_uint(w), C.c_uint(h), C.c_uint(frame), C.c_uint(bounces), C.c_uint(stride), C.c_uint(offset),
                acc.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p)]
        if self.kind == "port":
            args += [None, None]
        args.append(C.c_int(self.cores))
        t0 = time.perf_counter()
        getattr(self.lib, self.fn)(*args)
        return time.perf_counter() - t0


def cpu_sample(P, sc, w, h, bounces, steps, warmup, target_step_s, prefer_reference=True):
    """Times `steps` sampled launches; the stride is sized from one probe so a step takes about
    target_step_s on this box's cores."""
    cpu = CpuRenderer(prefer_reference)
    acc = np.zeros((h, w, 3), np.float32)
    out = np.zeros((h, w), np.uint32)
    n_blocks = (w // 16 + 1) * (h // 16 + 1)
    probe_stride = max(1, n_blocks // 256)
    t = cpu.run(sc, w, h, 1, bounces, probe_stride, 0, acc, out)
    full = t * probe_stride
    stride = int(max(1, min(n_blocks, round(full / target_step_s))))
    for k in range(warmup):
        cpu.run(sc, w, h, k + 1, bounces, stride, k % stride, acc, out)
    secs = segs = px = 0
    for k in range(steps):
        frame, off = warmup + k + 1, (warmup + k) % stride
        p, s = cpu.count(sc, w, h, frame, bounces, stride, off)
        secs += cpu.run(sc, w, h, frame, bounces, stride, off, acc, out)
        segs, px = segs + s, px + p
    return {"kind": cpu.kind, "cores": cpu.cores, "seconds": secs, "segments": segs, "pixels": px,
            "stride": stride, "mrays": segs / secs / 1e6, "steps": steps,
            "sample": f"{steps} launches x every {stride}-th 16x16 block of the {w}x{h} frame "
                      f"({px} px, {segs} segments, {secs:.1f} s on {cpu.cores} threads)"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    P = pkg()
    sc = load_scene(P, args.scene)
    w, h = (args.width, args.height) if args.width else (WEAK_DIMS.get(args.gpus, WEAK_DIMS[1])
                                                        if args.scaling == "weak" else WEAK_DIMS[1])
    total = max(1, args.steps + args.warmup)
    r = cpu_sample(P, sc, w, h, args.bounces, args.steps, args.warmup, min(2.0, 150.0 / total))
    ms = r["seconds"] / args.steps * 1e3
    frames_full = r["pixels"] / float(w * h)
    line = {"impl": "reference", "metric": "Mrays/s @1080p 1spp, README indoor scene", "value": r["mrays"], "unit": "Mrays/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "frames_per_s": frames_full / r["seconds"],
            "config": config_dict(args, w, h, sc),
            "sampling": "host cores; each step = a strided sample of 16x16 blocks of the frame",
            "cpu_baseline": {"value": r["mrays"], "unit": "Mrays/s", "cores": r["cores"], "kind": r["kind"],
                             "sample": r["sample"]},
            "e2e": {"value": r["mrays"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


WORKLOADS = {
    # name: (scene, width, height, bounces, scaling)
    "headline": ("indoor_tex", 0, 0, 3, "weak"),       # configs[1]; N > 1: N x the pixels (WEAK_DIMS)
    "cfg0": ("indoor_tex", 640, 360, 4, "strong"),
    "cfg1b4": ("indoor_tex", 1920, 1080, 4, "strong"),
    "cfg2": ("indoor_tex", 3840, 2160, 8, "strong"),   # 16 frames, gather after the last, bit-identity check
    "cfg3": ("synth2236", 1920, 1080, 4, "strong"),
}


def baseline_config_name(scene, w, h, bounces):
    """Which entry of BASELINE.json `configs` these arguments are (None: none of them)."""
    if scene.startswith("indoor"):
        if (w, h, bounces) == (640, 360, 4):
            return "configs[0] (README indoor scene 640x360, 1 spp, 4 bounces)"
        if (w, h) == (1920, 1080):
            return ("configs[1] (indoor scene 1920x1080, 1 spp/frame, temporal accumulation"
                    + ("" if bounces == 3 else f"; {bounces} iterations instead of the reference's constant 3") + ")")
        if (w, h, bounces) == (3840, 2160, 8):
            return "configs[2] (indoor scene 3840x2160, 8 bounces, image tiles across the GPUs)"
        if h * 16 == w * 9:
            return "configs[1] weak-scaled (the same 16:9 view at N x 1080p pixels)"
    if scene == "synth2236" and (w, h) == (1920, 1080):
        return "configs[3] (synthetic 10 M-triangle mesh, 1080p, glossy / refractive materials)"
    return None


def config_dict(args, w, h, sc):
    """Identical for both arms (the driver compares it)."""
    name = baseline_config_name(args.scene, w, h, args.bounces)
    return {"workload": f"{args.scene} {w}x{h} 1spp/frame bounces={args.bounces} temporal accumulation; "
                        + (f"BASELINE.json {name}" if name else "not a BASELINE.json configuration")
                        + ("; cubemap = the reference's constant fallback" if args.scene.startswith("indoor") else ""),
            "scene": args.scene, "width": w, "height": h, "bounces": args.bounces, "faces": sc.n_faces,
            "partition": (f"{args.gpus} ranks, round-robin 16-row strips, scene replicated"
                          if args.gpus > 1 and args.partition == "strips" else
                          f"{args.gpus} contiguous band(s) of whole 16-row strips, scene replicated")}


L2_POLICY = ("512 MB memset between timed frames (outside the per-frame event pairs); steady-state figure "
             "without flush alongside")


# ---------------------------------------------------------------------------------------------
# this repository's arm
# ---------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the path-tracing core has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    P = pkg()
    sc = load_scene(P, args.scene)
    n = max(world, 1)
    weak = args.scaling == "weak"
    w, h = (args.width, args.height) if args.width else (WEAK_DIMS.get(n, (1920 * n, 1080)) if weak else WEAK_DIMS[1])
    core = P.Core(local)
    sid = core.upload_scene(sc)
    cid = core.set_cubemap(sc.cubemap)
    # N > 1: round-robin 16-row strips (strip s -> rank s % N) so every rank sees the same mix
    # of cheap (sky / far wall) and expensive rows; --partition bands gives contiguous bands
    interleaved = world > 1 and args.partition == "strips"
    geom = P.tiles.BandGather(h, w, world, rank, interleaved=interleaved)
    strips = geom.strips
    r0, r1 = (0, 0) if interleaved else geom.band
    rows_kw = {"strips": strips} if interleaved else {"rows": (r0, r1)}
    bh = geom.n_rows
    dev = torch.device("cuda", local)
    accum = torch.zeros(max(bh, 1), w, 3, device=dev)
    out = torch.zeros(max(bh, 1), w, dtype=torch.int32, device=dev)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream()
    cs = stream.cuda_stream
    K, W = args.steps, args.warmup
    # N > 1, --gather store (default): every frame of every rank is stored by its kernels straight
    # into the full image on rank 0 (tiles.StoreAssembly; no collective), and rank 0's stream holds
    # the frame complete when all ranks' flags are in -- the image is assembled EVERY frame, inside
    # the per-frame timing.  --gather nccl: round 1's one all_gather_into_tensor after the last frame.
    store = P.tiles.StoreAssembly(P.core, local, h, w, world, rank) if world > 1 and args.gather == "store" else None
    # --gather copy (default): as above, but the pixels of ranks != 0 travel by DMA -- local band, pitched
    # peer copy on a copy stream overlapped with the next frame -- and rank 0 holds frame k complete
    # while it renders frame k + 1 (tiles.CopyAssembly); finish() closes the last frame inside the timing.
    if world > 1 and args.gather == "copy":
        if interleaved:
            store = P.tiles.CopyAssembly(P.core, local, h, w, world, rank, torch, out)
        else:   # contiguous bands: the strip scatter does not apply, the kernels store into the image
            store = P.tiles.StoreAssembly(P.core, local, h, w, world, rank)
    lagged = isinstance(store, P.tiles.CopyAssembly)
    seq = [0]  # frames issued through the assembly so far (flag values must only grow)

    def frame(k, host=None):
        if lagged:
            seq[0] += 1
            store.frame(lambda o, full, st: bh and core.render(sid, cid, sc, w, h, accum, o, None, bounces=args.bounces,
                                                               frame_nb=k, stream=st, out_full=full, **rows_kw),
                        strips, stream, host)
        elif store is not None:
            seq[0] += 1
            f = seq[0]
            store.before_frame(f, cs)
            if bh:
                core.render(sid, cid, sc, w, h, accum, store.image(f), None, bounces=args.bounces, frame_nb=k,
                            stream=cs, out_full=True, **rows_kw)
            store.after_frame(f, cs)
            if rank == 0:
                store.wait_complete(f, cs)
                if host is not None:        # e2e: the assembled image goes to pinned host memory
                    store.to_host(f, host, cs)
                store.consumed(f, cs)
        elif bh:
            core.render(sid, cid, sc, w, h, accum, out, None, bounces=args.bounces, frame_nb=k, stream=cs, **rows_kw)

    def barrier():
        if lagged:
            store.finish(stream)    # no frame left open (and no spinning wait left) across a synchronisation
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident value -----------------------------------------------------------
    for k in range(1, W + 1):
        frame(k)
    if world > 1 and store is None:
        geom.gather(out.view(torch.uint8).view(bh, w, 4))  # warm NCCL
    barrier()
    accum.zero_()   # the timed frames are samples 1..K of a fresh accumulation (config 2: compared bit for bit)
    core.reset_counters()
    launches_before = core.launch_count()
    sampler = ClockSampler(local) if rank == 0 else None
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_wall0 = time.time()
    for k in range(K):
        flush.zero_()                       # L2 flush between timed frames (outside the event pair)
        ev[k][0].record(stream)
        frame(1 + k)
        ev[k][1].record(stream)
    g0.record(stream)
    if world > 1 and store is None and bh:
        geom.gather(out.view(torch.uint8).view(bh, w, 4))
    if lagged:
        store.finish(stream)        # the last frame's assembly, inside the timed region
    g1.record(stream)
    barrier()
    t_wall1 = time.time()
    clocks = sampler.stop(t_wall0, t_wall1) if sampler else None
    per_frame = sorted(a.elapsed_time(b) for a, b in ev)
    ms_frames = sum(per_frame)
    ms_gather = g0.elapsed_time(g1)
    ms_total = ms_frames + ms_gather
    c = core.counters()
    launches = core.launch_count() - launches_before
    stats = torch.tensor([ms_total, float(c.segments), float(c.traced), float(launches), ms_gather],
                         dtype=torch.float64, device=dev)
    if world > 1:
        mx = stats.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)
        ms_total, ms_gather = float(mx[0]), float(mx[4])
    segments, traced, launches = float(stats[1]), float(stats[2]), int(stats[3])
    value = segments / (ms_total * 1e-3) / 1e6
    if os.environ.get("PTC_BENCH_DEBUG"):
        print(f"[rank {rank}] frames {ms_frames:.3f} ms  gather {ms_gather:.3f} ms  rows {bh}", file=sys.stderr, flush=True)

    # ---- N-GPU result == 1-GPU result, bit for bit (config 2's acceptance test) -------------
    # The band accumulators (samples 1..K) are gathered on every rank (torch.distributed, outside
    # any timing), rank 0 renders the same K samples of the WHOLE frame on its own GPU and compares
    # the float3 accumulator bit for bit; with --gather store also the assembled RGBA8 image.
    identity = None
    if args.verify:
        full_td = geom.gather(torch.flip(accum, dims=[0])[:bh].contiguous()) if bh or world > 1 else accum
        if rank == 0:
            acc1 = torch.zeros(h, w, 3, device=dev)
            out1 = torch.zeros(h, w, dtype=torch.int32, device=dev)
            ref_core = core
            for k in range(1, K + 1):
                ref_core.render(sid, cid, sc, w, h, acc1, out1, None, bounces=args.bounces, frame_nb=k, stream=cs)
            torch.cuda.synchronize()
            same_acc = bool(torch.equal(torch.flip(full_td, dims=[0]).view(torch.int32), acc1.view(torch.int32)))
            identity = {"accumulator_bit_identical_to_1gpu": same_acc, "samples": K, "pixels": w * h}
            if store is not None:
                img = np.zeros((h, w), np.uint32)
                store.to_host(seq[0], img, cs, sync=True)
                identity["image_bit_identical_to_1gpu"] = bool((img == out1.cpu().numpy().view(np.uint32)).all())
            del acc1, out1
        barrier()

    # ---- sustained: >= 5 s of back-to-back frames, no flush, clocks sampled ---------------------
    barrier()
    n_sus = max(K, int(np.ceil(args.sustain_s * 1e3 / max(ms_total / K, 1e-3)))) if args.sustain_s > 0 else 0
    sustained = None
    if n_sus:
        sus_sampler = ClockSampler(local) if rank == 0 else None
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        tw0 = time.time()
        s0.record(stream)
        for k in range(n_sus):
            frame(K + 1 + k)
        if lagged:
            store.finish(stream)
        s1.record(stream)
        barrier()
        tw1 = time.time()
        t = torch.tensor([s0.elapsed_time(s1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        sus_clocks = sus_sampler.stop(tw0, tw1) if sus_sampler else None
        sustained = {"frames": n_sus, "seconds": float(t[0]) * 1e-3, "ms_per_step": float(t[0]) / n_sus,
                     "value": segments / K * n_sus / (float(t[0]) * 1e-3) / 1e6, "unit": "Mrays/s",
                     "clocks": sus_clocks, "l2": "no flush between frames"}
    ms_steady = sustained["ms_per_step"] if sustained else None

    # ---- end to end through host buffers ----------------------------------------------------
    # N = 1: ptc_render_host_async -- arguments + camera host->device and the RGBA8 frame
    # device->pinned host EVERY step; the copy of frame k overlaps the rendering of frame k+1 (two
    # host buffers).  N > 1 (--gather store): every step ends with the ASSEMBLED full image in
    # rank 0's pinned host memory (the consumer the reference's window is, gpu_processor.cpp:365-386).
    e2e_d2h = int(bh) * w * 4
    e2e_what = "every rank's RGBA8 band -> its pinned host memory every step"
    barrier()
    if world > 1 and interleaved:
        # one page-locked host image in POSIX shared memory, mapped by every rank's process; each rank's
        # strips go straight to their rows over that rank's own PCIe link (ptc_memcpy_strips_to_host) on
        # a copy stream, two device bands / two host images alternating, so the copy of frame k overlaps
        # the rendering of frame k + 1.  The full image is assembled on the host every step without
        # crossing one GPU's link (66 MB per frame at N = 8 would take two frame times over one).
        box = [f"/ptc_bench_{os.getpid()}"]
        dist.broadcast_object_list(box, src=0)
        himg = P.core.HostImage(box[0], h, w, slots=2, create=True) if rank == 0 else None
        barrier()
        if rank != 0:
            himg = P.core.HostImage(box[0], h, w, slots=2, create=False)
        bands = [out, torch.zeros_like(out)]
        copy_stream = torch.cuda.Stream(device=dev)
        ev_r = [torch.cuda.Event() for _ in range(2)]
        ev_c = [torch.cuda.Event() for _ in range(2)]

        def host_frame(k, i):
            b = i & 1
            if bh:
                stream.wait_event(ev_c[b])      # the copy that last read this device band is done
                core.render(sid, cid, sc, w, h, accum, bands[b], None, bounces=args.bounces, frame_nb=k, stream=cs,
                            **rows_kw)
                ev_r[b].record(stream)
                copy_stream.wait_event(ev_r[b])
                himg.put_strips(local, b, bands[b], strips, stream=copy_stream.cuda_stream)
                ev_c[b].record(copy_stream)

        for k in range(W):
            host_frame(k + 1, k)
        copy_stream.synchronize()
        barrier()
        t0 = time.perf_counter()
        for k in range(K):
            host_frame(W + 1 + k, k)
        copy_stream.synchronize()
        torch.cuda.synchronize()
        e2e_what = ("every rank's strips -> their rows of ONE page-locked host image in shared memory (assembled on the "
                    "host every step, each rank over its own PCIe link, copy overlapped with the next frame)")
    elif store is not None:
        h_img = [torch.zeros(h, w, dtype=torch.int32).pin_memory() for _ in range(2)] if rank == 0 else None
        for k in range(W):
            frame(k + 1, h_img[k & 1].numpy() if rank == 0 else None)
        barrier()
        t0 = time.perf_counter()
        for k in range(K):
            frame(W + 1 + k, h_img[k & 1].numpy() if rank == 0 else None)
        torch.cuda.synchronize()
        e2e_d2h = w * h * 4 if rank == 0 else 0
        e2e_what = "assembled full image -> rank 0 pinned host memory every step"
    else:
        h_accum = torch.zeros(max(bh, 1), w, 3).pin_memory()
        h_out = [torch.zeros(max(bh, 1), w, dtype=torch.int32).pin_memory() for _ in range(2)]
        if bh:
            for k in range(1, W + 1):
                core.render_host_async(sid, cid, sc, w, h, h_accum.numpy(), h_out[k & 1].numpy(), None,
                                       bounces=args.bounces, frame_nb=k, upload_accum=(k == 1), stream=cs, **rows_kw)
            core.wait()
        barrier()
        t0 = time.perf_counter()
        for k in range(K):
            if bh:
                core.render_host_async(sid, cid, sc, w, h, h_accum.numpy(), h_out[k & 1].numpy(), None,
                                       bounces=args.bounces, frame_nb=W + 1 + k, stream=cs, **rows_kw)
        if bh:
            core.wait()
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3  # host clock: the call chain ends with frames on the host
    t = torch.tensor([e2e_ms, float(e2e_d2h)], dtype=torch.float64, device=dev)
    if world > 1:
        mx = t.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        e2e_ms = float(mx[0])
    e2e_d2h_total = int(t[1])
    e2e_value = segments / (e2e_ms * 1e-3) / 1e6

    # ---- roofline of the stage kernels, baselines (rank 0) ----------------------------------
    roof = cpu_base = gpu_base = None
    if rank == 0:
        sm_mhz = (clocks or {}).get("sm_mhz")
        roof = roofline(P, core, sid, cid, sc, w, h, rows_kw, args, accum, out, stream, sm_mhz=sm_mhz)
        if world == 1 and not args.no_cpu_baseline:
            try:
                r = cpu_sample(P, sc, w, h, args.bounces, 3, 1, 4.0, prefer_reference=False)
                cpu_base = {"value": r["mrays"], "unit": "Mrays/s", "cores": r["cores"], "kind": r["kind"],
                            "sample": r["sample"]}
            except Exception as e:  # the checker is optional for the product, mandatory for the report
                cpu_base = {"value": None, "unit": "Mrays/s", "cores": 0, "kind": "port", "sample": f"failed: {e}"}
            if sc.n_faces <= 100_000:
                gpu_base = reference_kernel_on_gpu(sc, w, h, args.bounces, K, W)
            else:
                gpu_base = {"unavailable": f"brute force over {sc.n_faces} faces per ray: not runnable at this size"}

    if rank == 0:
        line = {"metric": "Mrays/s @1080p 1spp, README indoor scene", "value": value, "unit": "Mrays/s", "n_gpus": world,
                "steps": K, "warmup": W, "ms_per_step": ms_total / K, "higher_is_better": True,
                "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "frames_per_s": K / (ms_total * 1e-3), "segments_per_step": segments / K,
                "traced_per_step": traced / K, "ms_per_step_steady_no_flush": ms_steady,
                "ms_gather_once": ms_gather,
                "gather": ("none (1 GPU)" if world == 1 else
                           "every frame: each rank's strips DMA-copied (pitched peer copy over NVLink, CUDA IPC) into the "
                           "full image on rank 0, overlapped with the next frame; rank 0's stream waits for all ranks' "
                           "flags of frame k while frame k + 1 renders, the last frame's wait is inside the timed region "
                           "(ms_gather_once); no collective" if lagged else
                           "every frame: kernels store RGBA8 into the full image on rank 0 over NVLink (CUDA IPC), "
                           "rank 0's stream waits for all ranks' flags inside the per-frame timing; no collective"
                           if store is not None else "one NCCL all_gather_into_tensor after the last timed frame"),
                "ms_per_frame_rank0": {"min": per_frame[0], "median": per_frame[len(per_frame) // 2],
                                       "max": per_frame[-1]},
                "config": config_dict(args, w, h, sc), "l2_policy": L2_POLICY,
                "clocks": clocks, "sustained": sustained, "bit_identity": identity,
                "e2e": {"value": e2e_value, "unit": "Mrays/s", "ms_per_step": e2e_ms / K,
                        "h2d_bytes_per_step": (C.sizeof(P.core.PtcRenderArgs) + 64) * world,
                        "d2h_bytes_per_step": e2e_d2h_total,
                        "what": e2e_what},
                "gpu_launches": launches, "roofline": roof, "cpu_baseline": cpu_base,
                "reference_kernel_same_gpu": gpu_base}
        print(json.dumps(line))
    barrier()
    if world > 1 and interleaved:
        himg.close()
    if store is not None:
        store.close()
    core.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def ncu_counters(args, w, h):
    """Per-frame counters of this workload's stage kernels from the committed ncu capture
    (profiles/r2_counters.json, written by tools/ncu_counters.py from `ncu --csv` of tools/probe.py):
    warp instructions executed and DRAM bytes.  NOT measured in this run -- the entry says which
    capture it came from; None when the workload has no capture."""
    try:
        table = json.load(open(os.path.join(REPO, "profiles", "r2_counters.json")))
        return table.get(f"{args.scene} {w}x{h} b={args.bounces}")
    except Exception:
        return None


def roofline(P, core, sid, cid, sc, w, h, rows_kw, args, accum, out, stream, sm_mhz=None):
    """Three lower bounds on the time of a frame's stage kernels (k_path / k_extend; every launch of
    a frame is one of them), all per frame:
      bytes  BVH bytes (nodes visited x 128 B + triangle records tested x 48 B, counted by the
             kernels' statistics pass, which tests pin to oracle O2's counts) at the MEASURED read
             bandwidth of the cache tier the BVH lives in + every other algorithmic byte (96 B of
             shading record and texel per traced ray, 2 x 64 B per queued ray, 28 B per pixel) at
             the HBM peak of MEASURED_PEAKS.json
      fp32   nodes x 160 + triangles x 37 flops at SMs x 128 lanes x 2 x the SM clock
      issue  warp instructions (committed ncu capture) at SMs x 4 schedulers x the SM clock
    `bound` = the largest, `frac` = that bound / the kernels' CUDA-event time (events around every
    launch, one pipeline).  The algorithmic-bytes-over-HBM-peak figure of round 1 is kept under
    `algorithmic_gbs_vs_hbm` -- it is not a roofline fraction for a cache-resident scene."""
    hbm, how = measured_peaks()
    try:
        prof = core.profile_frame(sid, cid, sc, w, h, accum, out, None, bounces=args.bounces, frame_nb=5,
                                  stream=stream.cuda_stream, repeats=8, **rows_kw)
    except Exception as e:
        return {"bound": None, "achieved": None, "peak": None, "unit": None, "frac": None, "traffic": None,
                "note": f"profile_frame failed: {e}"}
    info = core.scene_info(sid)
    px = core.band_rows(w, h, **rows_kw) * w
    b_bvh = prof["nodes_visited"] * info.bvh_node_bytes + prof["tris_tested"] * info.tri_bytes
    b_shade = prof["traced"] * (info.shade_bytes + 16)
    b_queue = prof["queue_rays"] * (prof["queue_bytes_per_ray"] * 2 + prof["hit_bytes_per_ray"] * 2
                                    + (32 if prof["hit_bytes_per_ray"] else 0))
    b_fb = px * 28
    nbytes = b_bvh + b_shade + b_queue + b_fb
    k_ms = sum(prof["launch_ms"])
    n = max(1, len(prof["launch_ms"]))
    clock_hz = (sm_mhz or 1965.0) * 1e6
    n_sms = 148
    flops = prof["nodes_visited"] * 160.0 + prof["tris_tested"] * 37.0
    fp32_peak = n_sms * 128 * 2 * clock_hz
    issue_peak = n_sms * 4 * clock_hz
    bvh_resident = info.bvh_nodes * info.bvh_node_bytes + info.n_faces * info.tri_bytes
    tier = "l1" if bvh_resident <= (64 << 10) else ("l2" if bvh_resident <= (64 << 20) else "hbm")
    tiers = {"l1_gbs": core.read_bandwidth(32 << 10, 48 << 30), "l2_gbs": core.read_bandwidth(32 << 20, 24 << 30),
             "hbm_read_gbs": core.read_bandwidth(4 << 30, 16 << 30), "bvh_resident_bytes": bvh_resident,
             "bvh_tier": tier,
             "how": "16-byte read-only loads from all SMs over a zeroed working set of 32 KB / 32 MB / 4 GB"}
    tier_bw = {"l1": tiers["l1_gbs"], "l2": tiers["l2_gbs"], "hbm": min(tiers["hbm_read_gbs"], hbm)}[tier]
    ctr = ncu_counters(args, w, h)
    bytes_model = "BVH bytes at the measured bandwidth of their cache tier + the rest at the HBM peak"
    if tier == "hbm" and ctr:
        # a BVH beyond L2: its upper levels are still cache hits, so charging every algorithmic byte to
        # HBM would overstate the bound (round-1 review).  The bytes that DID come from DRAM (committed
        # ncu capture of this workload) are charged at the HBM peak, the remainder at the measured L2
        # bandwidth.
        dram = min(float(ctr["dram_bytes_per_frame"]), float(nbytes))
        t_bytes = (dram / hbm + (nbytes - dram) / tiers["l2_gbs"]) / 1e9 * 1e3
        bytes_model = ("measured DRAM bytes (committed ncu capture) at the HBM peak + the remaining algorithmic bytes "
                       "at the measured L2 bandwidth")
    else:
        t_bytes = (b_bvh / tier_bw + (nbytes - b_bvh) / hbm) / 1e9 * 1e3
    t_fp32 = flops / fp32_peak * 1e3
    t_issue = ctr["warp_instructions_per_frame"] / issue_peak * 1e3 if ctr else None
    bounds = {"bytes": t_bytes, "fp32": t_fp32}
    if t_issue is not None:
        bounds["issue"] = t_issue
    bound = max(bounds, key=bounds.get)
    if bound == "issue":
        ach, peak, unit = ctr["warp_instructions_per_frame"] / (k_ms * 1e-3) / 1e9, issue_peak / 1e9, "Gwarp-inst/s"
    elif bound == "fp32":
        ach, peak, unit = flops / (k_ms * 1e-3) / 1e12, fp32_peak / 1e12, "TFLOP/s"
    else:
        ach, peak, unit = nbytes / (k_ms * 1e-3) / 1e9, nbytes / (t_bytes * 1e-3) / 1e9, "GB/s"
    traffic = ctr["dram_bytes_per_frame"] / n if ctr else None
    return {"bound": {"bytes": tier if tier != "hbm" else "hbm", "fp32": "fp32", "issue": "issue"}[bound],
            "achieved": ach, "peak": peak, "unit": unit, "frac": bounds[bound] / k_ms, "traffic": traffic,
            "traffic_source": (ctr["source"] + " (committed capture, not measured in this run)") if ctr else None,
            "traffic_over_algorithmic": (ctr["dram_bytes_per_frame"] / nbytes) if ctr else None,
            "bounds_ms": bounds, "bounds_frac": {k: v / k_ms for k, v in bounds.items()},
            "sm_clock_mhz_used": clock_hz / 1e6,
            "peak_note": "bytes: effective peak = algorithmic bytes / (" + bytes_model + "); hbm peak " + how,
            "cache_tiers": tiers, "hbm_peak_gbs": hbm,
            "algorithmic_gbs_vs_hbm": {"achieved": nbytes / (k_ms * 1e-3) / 1e9, "peak": hbm,
                                       "ratio": nbytes / (k_ms * 1e-3) / 1e9 / hbm,
                                       "note": "NOT a roofline fraction when the BVH is cache-resident"},
            "kernel": "ptc::k_path + ptc::k_extend (the wavefront stages of one frame)",
            "launches_per_frame": n, "kernel_ms_per_launch": k_ms / n, "kernel_ms_per_frame": k_ms,
            "frame_ms_one_pipeline": prof["frame_ms"], "frame_ms_as_shipped": prof["frame_ms_shipped"],
            "pipelines_as_shipped": prof["pipelines"],
            "frac_vs_shipped_frame": bounds[bound] / prof["frame_ms_shipped"] if prof["frame_ms_shipped"] else None,
            "kernel_share_of_frame": k_ms / prof["frame_ms"],
            "launch_ms": prof["launch_ms"], "launch_kind": prof["launch_kind"],
            "algorithmic_bytes_per_launch": nbytes / n,
            "algorithmic_bytes_per_frame": {"bvh": b_bvh, "shade": b_shade, "queue": b_queue, "framebuffer": b_fb},
            "intersection_flops_per_frame": flops,
            "warp_instructions_per_frame": ctr["warp_instructions_per_frame"] if ctr else None,
            "nodes_per_ray": prof["nodes_visited"] / max(1, prof["traced"]),
            "tris_per_ray": prof["tris_tested"] / max(1, prof["traced"]),
            "extend_lanes_per_warp_iteration": ([x / max(1, prof["extend_lanes"][0]) for x in prof["extend_lanes"][1:]]
                                                if prof["extend_lanes"][0] else None),
            "rays_per_bounce": prof["rays_per_bounce"][:args.bounces]}


def reference_kernel_on_gpu(sc, w, h, bounces, K, W):
    """GPU baseline the north star names: the reference's own kernel (CUDA-12 patch, sm_100a,
    oracle/_ref/libref_gpu.so) timed on the same B200 with the same frames."""
    import torch
    path = os.path.join(REPO, "oracle", "_ref", "libref_gpu.so")
    if not os.path.exists(path):
        return {"unavailable": "oracle/_ref/libref_gpu.so not built (needs /root/reference at build time)"}
    sys.path.insert(0, os.path.join(REPO, "tests"))
    from pt_testlib import RefGpu
    g0 = RefGpu(sc)
    acc = torch.zeros(h, w, 3, device="cuda")
    out = torch.zeros(h, w, dtype=torch.int32, device="cuda")
    for k in range(1, W + 1):
        g0.kernel(w, h, k, bounces, acc.data_ptr(), out.data_ptr())
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    K = min(K, 32)
    e0.record()
    for k in range(K):
        g0.kernel(w, h, W + 1 + k, bounces, acc.data_ptr(), out.data_ptr())
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / K
    g0.close()
    return {"ms_per_step": ms, "value": w * h * bounces / (ms * 1e-3) / 1e6, "unit": "Mrays/s",
            "note": "reference raytrace.cu kernel, brute force over all faces, 16x16 blocks, same seeds"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scene", default="indoor_tex")
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--bounces", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--partition", default="strips", choices=["strips", "bands"],
                    help="N > 1: round-robin 16-row strips (default) or contiguous bands")
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS),
                    help="a BASELINE.json configuration (sets scene / size / bounces / scaling)")
    ap.add_argument("--scaling", default=None, choices=["weak", "strong"])
    ap.add_argument("--gather", default="copy", choices=["copy", "store", "nccl"],
                    help="N > 1: image assembled on rank 0 every frame by overlapped peer DMA copies (default) or by the "
                         "kernels' own peer stores, or one NCCL gather at the end")
    ap.add_argument("--verify", action="store_true",
                    help="compare the N-GPU accumulator (and image) bit for bit with a 1-GPU render on rank 0")
    ap.add_argument("--sustain-s", type=float, default=5.0, help="length of the sustained block (0 = skip)")
    args = ap.parse_args()
    if args.workload:
        scene, ww, hh, bb, scaling = WORKLOADS[args.workload]
        args.scene, args.bounces = scene, bb
        if ww:
            args.width, args.height = ww, hh
        args.scaling = args.scaling or scaling
        if args.workload == "cfg2":
            args.verify = True
    if not args.scaling:
        args.scaling = "strong" if args.width else "weak"
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
