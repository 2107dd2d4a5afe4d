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


class BandGather:
    """Gathers per-rank bands into the full image with one all_gather.

    ``gather(t)``: ``t`` is this rank's band, shape (band_rows, width, ...) in top-down row
    order (RGBA8 out buffer as is; accumulator via :func:`accum_band_to_image_rows` or
    ``flip_rows=True``).  Bands are padded to the largest band so a single equal-size
    ``all_gather_into_tensor`` (ncclAllGather on the box) moves everything; the result is the
    full image on every rank.
    """

    def __init__(self, height: int, width: int, world: int, rank: int, strip: int = STRIP):
        self.height, self.width, self.world, self.rank = height, width, world, rank
        self.bands = row_bands(height, world, strip)
        self.max_rows = max(e - b for b, e in self.bands)
        self._pad = {}
        self._all = {}

    @property
    def band(self) -> Tuple[int, int]:
        return self.bands[self.rank]

    def gather(self, t, flip_rows: bool = False):
        import torch
        import torch.distributed as dist

        b, e = self.band
        if t.shape[0] != e - b or t.shape[1] != self.width:
            raise ValueError(f"band tensor has shape {tuple(t.shape)}, expected ({e - b}, {self.width}, ...)")
        if flip_rows:
            t = torch.flip(t, dims=[0])
        key = (t.dtype, t.device, tuple(t.shape[2:]))
        if key not in self._pad:
            self._pad[key] = torch.zeros((self.max_rows, self.width) + tuple(t.shape[2:]),
                                         dtype=t.dtype, device=t.device)
            self._all[key] = torch.empty((self.world * self.max_rows, self.width) + tuple(t.shape[2:]),
                                         dtype=t.dtype, device=t.device)
        pad, out = self._pad[key], self._all[key]
        pad[: e - b].copy_(t)
        if self.world == 1:
            out.copy_(pad)
        else:
            dist.all_gather_into_tensor(out, pad)
        if all(eb - bb == self.max_rows for bb, eb in self.bands):
            return out[: self.height]
        parts = [out[r * self.max_rows: r * self.max_rows + (eb - bb)] for r, (bb, eb) in enumerate(self.bands)]
        return torch.cat(parts, dim=0)
