"""ptc_render_static -- the render call that allocates nothing -- against ptc_render.

Collected last on purpose.  Runs alone in a fresh process or
after the other GPU tests: it uploads its own scenes and uses its own buffers.  Both entry points
run the same kernels on the same device, so the bar is the one test_gpu_parity.py uses for such
pairs: accumulator, radiance and RGBA8 BIT-IDENTICAL.
"""
import numpy as np
import pytest

from pt_testlib import cube_of, load_scene, pkg

pytestmark = pytest.mark.gpu
P = pkg()


def bits(t):
    return np.ascontiguousarray(t.cpu().numpy()).view(np.uint32)


def dev_buffers(w, h):
    import torch
    return (torch.zeros(h, w, 3, device="cuda"), torch.zeros(h, w, dtype=torch.int32, device="cuda"),
            torch.zeros(h, w, 3, device="cuda"))


def upload(core, sc):
    if sc.cubemap is None:
        sc.cubemap = P.synth.gradient_cubemap(32)
        sc.cubemap_size = 32
    return core.upload_scene(sc), core.set_cubemap(cube_of(sc)[0])


def pair(core, sid, cid, sc, w, h, bounces, frames, **kw):
    """`frames` consecutive calls of each entry point into its own buffers; compares after each."""
    import torch
    a_s, o_s, r_s = dev_buffers(w, h)
    a_d, o_d, r_d = dev_buffers(w, h)
    for f in frames:
        core.render_static(sid, cid, sc, w, h, a_s, o_s, r_s, bounces=bounces, frame_nb=f, **kw)
        torch.cuda.synchronize()
        core.render(sid, cid, sc, w, h, a_d, o_d, r_d, bounces=bounces, frame_nb=f, **kw)
        torch.cuda.synchronize()
        assert a_s.any()
        assert (bits(a_s) == bits(a_d)).all(), f"accumulator differs at frame {f}"
        assert (bits(r_s) == bits(r_d)).all(), f"radiance differs at frame {f}"
        assert (o_s.cpu().numpy() == o_d.cpu().numpy()).all(), f"RGBA8 differs at frame {f}"


def test_static_scratch_render_equals_ptc_render(core):
    import torch
    sc = load_scene("indoor_tex")
    sid, cid = upload(core, sc)
    # two calls in a row (the second accumulates on the first), at the largest 16:9 size that fits
    pair(core, sid, cid, sc, 640, 360, 4, [1, 2])
    # a smaller frame right after a larger one: the array still holds the larger call's queues
    pair(core, sid, cid, sc, 250, 130, 3, [5, 6])
    # a scene whose schedule is the split form (extend kernel + hit records), 8 iterations
    big = P.synth.terrain_room(grid=40, tex_size=16)
    bsid, bcid = upload(core, big)
    pair(core, bsid, bcid, big, 320, 180, 8, [1, 2])
    # a band of a larger image, and the preview (moved) frame
    pair(core, sid, cid, sc, 1920, 1080, 3, [3, 4], rows=(400, 528))
    pair(core, sid, cid, sc, 320, 180, 3, [1], moved=True)
    torch.cuda.synchronize()


def test_static_scratch_render_allocates_nothing_and_rejects_large_frames(core):
    import torch
    sc = load_scene("indoor_plain")
    sid, cid = upload(core, sc)
    a, o, _ = dev_buffers(64, 36)
    core.render_static(sid, cid, sc, 64, 36, a, o, None, bounces=3, frame_nb=1)  # kernels loaded
    torch.cuda.synchronize()
    # a size this context has never rendered: ptc_render would allocate 176 B x 230 400 = 40 MB
    w, h = 640, 360
    a, o, _ = dev_buffers(w, h)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for f in (1, 2):
        core.render_static(sid, cid, sc, w, h, a, o, None, bounces=3, frame_nb=f)
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    assert free0 - free1 < (8 << 20), f"device memory shrank by {free0 - free1} bytes across the calls"
    assert a.any()

    # beyond the stated maximum: refused with a clear message, frame counter untouched
    assert w * h <= P.core.STATIC_MAX_PIXELS < 1920 * 1080
    A, O, _ = dev_buffers(1920, 1080)
    core.reset_accumulation()
    with pytest.raises(P.core.PtcError, match="static scratch limit"):
        core.render_static(sid, cid, sc, 1920, 1080, A, O, None, bounces=3)
    assert not A.any()
    # frame_nb = 0 follows the context's counter: the refused call must not have advanced it
    a1, o1, _ = dev_buffers(160, 90)
    a2, o2, _ = dev_buffers(160, 90)
    core.render_static(sid, cid, sc, 160, 90, a1, o1, None, bounces=3)          # counter -> 1
    core.render(sid, cid, sc, 160, 90, a2, o2, None, bounces=3, frame_nb=1)
    torch.cuda.synchronize()
    assert (bits(a1) == bits(a2)).all()
