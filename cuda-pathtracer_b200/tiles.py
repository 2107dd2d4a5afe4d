"""Image-tile partitioning across GPUs and the framebuffer gather.

The reference is single-GPU (cuda_opengl/src/driver/gpu_info.cpp:26-28,38-43).  Pixels are
independent units -- the RNG seed of a pixel depends only on its GLOBAL position
(raytrace.cu:227-235) -- so the image is split into bands of whole 16-row strips (the
reference's 16x16 blocks stay whole), the scene is replicated, every rank accumulates its
own band across frames, and ONE collective per reported image gathers the RGBA8 (or the
float3 accumulator) bands.  No collective sits on the per-frame path.

One process per GPU; ``torch.distributed`` (NCCL over NVLink on the box, gloo in CPU
tests) is the plumbing.  Python here is host logic only.
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np

STRIP = 16  # rows per strip == the reference's block height (raytrace.cu:315)


def row_bands(height: int, world: int, strip: int = STRIP) -> List[Tuple[int, int]]:
    """Contiguous row bands [begin, end), one per rank, each a whole number of strips
    (except that the last band ends at ``height``); bands differ by at most one strip.
    Ranks beyond the number of strips get an empty band (begin == end)."""
    if height <= 0 or world <= 0:
        raise ValueError("height and world must be positive")
    n_strips = (height + strip - 1) // strip
    base, extra = divmod(n_strips, world)
    out, s = [], 0
    for r in range(world):
        k = base + (1 if r < extra else 0)
        b, e = min(s * strip, height), min((s + k) * strip, height)
        out.append((b, e))
        s += k
    return out


def band_rows_max(height: int, world: int, strip: int = STRIP) -> int:
    return max(e - b for b, e in row_bands(height, world, strip))


def accum_band_to_image_rows(band_accum: np.ndarray) -> np.ndarray:
    """A band accumulator is stored bottom-up inside the band (raytrace.cu:252 applied to the
    band); returns it top-down so bands can be stacked in image order."""
    return band_accum[::-1]


def strip_rows_of(height: int, world: int, rank: int, strip: int = STRIP) -> np.ndarray:
    """Global rows of rank ``rank`` under round-robin strip interleaving (strip s goes to
    rank s % world), in band-local order -- the order ptc_render uses for
    ``strip_rows=strip, strip_stride=world, strip_phase=rank`` (device_types.cuh global_row)."""
    if height <= 0 or world <= 0 or not 0 <= rank < world:
        raise ValueError("bad height / world / rank")
    n_strips = (height + strip - 1) // strip
    parts = [np.arange(s * strip, min((s + 1) * strip, height)) for s in range(rank, n_strips, world)]
    return np.concatenate(parts) if parts else np.zeros(0, dtype=np.int64)


class BandGather:
    """Gathers per-rank bands into the full image with one all_gather.

    ``interleaved=False``: rank r owns the contiguous band ``row_bands(...)[r]``.
    ``interleaved=True``: rank r owns strips r, r+world, ... (``strip_rows_of``), which evens
    out scene-dependent cost between ranks; pass ``strips`` to ``Core.render``.

    ``gather(t)``: ``t`` is this rank's band, shape (band_rows, width, ...) in band-local
    top-down row order (RGBA8 out buffer as is; accumulator with ``flip_rows=True``).  Bands
    are padded to the largest band so a single equal-size ``all_gather_into_tensor``
    (ncclAllGather on the box) moves everything; one row gather then puts every row at its
    image position.  The result is the full image on every rank.
    """

    def __init__(self, height: int, width: int, world: int, rank: int, strip: int = STRIP,
                 interleaved: bool = False):
        self.height, self.width, self.world, self.rank = height, width, world, rank
        self.strip, self.interleaved = strip, interleaved
        self.bands = row_bands(height, world, strip)
        if interleaved:
            self.rows = [strip_rows_of(height, world, r, strip) for r in range(world)]
        else:
            self.rows = [np.arange(b, e) for b, e in self.bands]
        self.max_rows = max(len(r) for r in self.rows)
        # image row -> row of the gathered (world * max_rows) buffer
        src = np.zeros(height, dtype=np.int64)
        for r, rows in enumerate(self.rows):
            src[rows] = r * self.max_rows + np.arange(len(rows))
        self._src = src
        self._identity = bool(np.array_equal(src, np.arange(height)))
        self._pad = {}
        self._all = {}
        self._idx = {}

    @property
    def band(self) -> Tuple[int, int]:
        """Contiguous mode: [begin, end) of this rank."""
        if self.interleaved:
            raise ValueError("interleaved partition has no contiguous band; use .strips / .n_rows")
        return self.bands[self.rank]

    @property
    def n_rows(self) -> int:
        return len(self.rows[self.rank])

    @property
    def strips(self):
        """(strip_rows, strip_stride, strip_phase) for Core.render, or None when contiguous."""
        return (self.strip, self.world, self.rank) if self.interleaved else None

    def gather(self, t, flip_rows: bool = False):
        import torch
        import torch.distributed as dist

        n = self.n_rows
        if t.shape[0] != n or t.shape[1] != self.width:
            raise ValueError(f"band tensor has shape {tuple(t.shape)}, expected ({n}, {self.width}, ...)")
        if flip_rows:
            t = torch.flip(t, dims=[0])
        key = (t.dtype, t.device, tuple(t.shape[2:]))
        if key not in self._pad:
            self._pad[key] = torch.zeros((self.max_rows, self.width) + tuple(t.shape[2:]),
                                         dtype=t.dtype, device=t.device)
            self._all[key] = torch.empty((self.world * self.max_rows, self.width) + tuple(t.shape[2:]),
                                         dtype=t.dtype, device=t.device)
            self._idx[key] = torch.from_numpy(self._src).to(t.device)
        pad, out = self._pad[key], self._all[key]
        pad[:n].copy_(t)
        if self.world == 1:
            out.copy_(pad)
        else:
            dist.all_gather_into_tensor(out, pad)
        if self._identity:
            return out[: self.height]
        return out.index_select(0, self._idx[key])


class StoreAssembly:
    """One process per GPU: the image is assembled on rank 0 by the STORES of every rank's last
    stage kernel (PTC_OUT_FULL_IMAGE into rank 0's memory, mapped through CUDA IPC over NVLink) --
    no gather pass, no pad copy, no permutation, no collective on the data path.  torch.distributed
    only carries the 64-byte IPC handle at set-up.

    Two full-size images alternate (frame k -> image k & 1) so that rank 0 can hand frame k to its
    consumer while frame k + 1 is being stored.  Flags (one 128-byte line each, in rank 0's memory):
      ready[r]    = last frame whose strips rank r has stored        (written by rank r)
      consumed    = last frame rank 0's consumer is done with        (written by rank 0)
    Rank r renders frame k only after consumed >= k - 2 (the image it is about to overwrite is
    free); rank 0's stream holds frame k complete after ready[r] >= k for all r.
    """
    LINE = 128

    def __init__(self, core_mod, device, height, width, world, rank):
        import torch.distributed as dist
        self.m, self.device, self.world, self.rank = core_mod, device, world, rank
        self.h, self.w = height, width
        self.image_bytes = ((height * width * 4 + 255) // 256) * 256
        total = 2 * self.image_bytes + (world + 1) * self.LINE
        box = [None]
        if rank == 0:
            self.buf = core_mod.SharedBuffer(device, total)
            box[0] = self.buf.handle
        if world > 1:
            dist.broadcast_object_list(box, src=0)
        if rank != 0:
            self.buf = core_mod.SharedBuffer(device, handle=box[0])
        self.flags = self.buf.address + 2 * self.image_bytes

    def image(self, frame):
        return self.buf.address + (frame & 1) * self.image_bytes

    def before_frame(self, frame, stream):
        """Back-pressure: the image of `frame` was last used by frame - 2."""
        if self.rank != 0 and frame > 2:
            self.m.flag_wait(self.device, self.flags + self.world * self.LINE, 1, 1, frame - 2, stream)

    def after_frame(self, frame, stream):
        """This rank's strips of `frame` are stored (release at system scope, stream order)."""
        self.m.flag_signal(self.device, self.flags + self.rank * self.LINE, frame, stream)

    def wait_complete(self, frame, stream):
        """Rank 0: `stream` continues when every rank's strips of `frame` have landed."""
        assert self.rank == 0
        self.m.flag_wait(self.device, self.flags, self.world, self.LINE // 4, frame, stream)

    def consumed(self, frame, stream):
        """Rank 0: everything enqueued on `stream` so far is done with the image of `frame`."""
        assert self.rank == 0
        self.m.flag_signal(self.device, self.flags + self.world * self.LINE, frame, stream)

    def to_host(self, frame, out, stream, sync=False):
        assert self.rank == 0
        self.buf.to_host(out, offset=(frame & 1) * self.image_bytes, stream=stream, sync=sync)

    def close(self):
        self.buf.close()


class CopyAssembly(StoreAssembly):
    """Same image, same flags, but the pixels of ranks != 0 travel by DMA instead of by the kernels'
    stores: a rank renders into a local band, and a copy stream scatters the band's strips into rank
    0's image (ptc_memcpy_strips_to_device: one pitched peer copy over NVLink), then raises the
    rank's flag.  Two local bands alternate, so the copy of frame k overlaps the rendering of frame
    k + 1; rank 0 (whose own strips are stored by its kernels, locally) holds frame k complete while
    it renders frame k + 1 -- the consumer runs ONE frame behind, ``finish`` closes the last one.

    Why: seven GPUs storing 32-byte pixel runs into one GPU during the short last stage of a frame
    cost 0.16 ms per 0.62 ms frame on 8 B200 (StoreAssembly: 0.778 ms/frame); see DESIGN.md 6.
    """

    def __init__(self, core_mod, device, height, width, world, rank, torch, band_like):
        super().__init__(core_mod, device, height, width, world, rank)
        self.torch = torch
        self.frames = 0      # frames issued
        self.closed = 0      # frames whose completion rank 0 has waited for
        if rank != 0:
            self.copy_stream = torch.cuda.Stream(device=band_like.device)
            self.bands = [torch.zeros_like(band_like), torch.zeros_like(band_like)]
            self.ev_r = [torch.cuda.Event() for _ in range(2)]
            self.ev_c = [torch.cuda.Event() for _ in range(2)]

    def frame(self, render, strips, stream, host=None):
        """One frame.  ``render(out, out_full, stream_handle)`` enqueues this rank's strips;
        ``stream`` is the torch stream of the render.  Rank 0: ``host`` (numpy) receives the frame
        that completes during this call (the previous one)."""
        self.frames += 1
        f, cs = self.frames, stream.cuda_stream
        if self.rank == 0:
            render(self.image(f), True, cs)          # own strips: local stores into the image
            self.after_frame(f, cs)
            if f > 1:
                self._close(f - 1, cs, host)
            return
        b = f & 1
        stream.wait_event(self.ev_c[b])              # the copy that last read this band is done
        render(self.bands[b], False, cs)
        self.ev_r[b].record(stream)
        self.copy_stream.wait_event(self.ev_r[b])
        ccs = self.copy_stream.cuda_stream
        if f > 2:                                     # the image of frame f was last used by frame f - 2
            self.m.flag_wait(self.device, self.flags + self.world * self.LINE, 1, 1, f - 2, ccs)
        self.m.strips_to_device(self.device, self.image(f), self.bands[b], self.w, self.h, strips, ccs)
        self.m.flag_signal(self.device, self.flags + self.rank * self.LINE, f, ccs)
        self.ev_c[b].record(self.copy_stream)

    def _close(self, f, cs, host=None):
        self.wait_complete(f, cs)
        if host is not None:
            self.to_host(f, host, cs)
        self.consumed(f, cs)
        self.closed = f

    def finish(self, stream, host=None):
        """Rank 0: wait for the last issued frame too (inside the caller's timing)."""
        if self.rank == 0 and self.closed < self.frames:
            self._close(self.frames, stream.cuda_stream, host)

