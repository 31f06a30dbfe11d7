"""Multi-GPU partition of the pass: contiguous row bands of the final level, one per rank.

A pixel of level i depends only on its own ancestor chain (x>>k, y>>k) of the coarser levels
(compute.glsl:81) and on uniforms, so ranks never exchange data while tracing: each rank traces its
band plus the few coarse rows its band's ancestors touch (recomputed redundantly, bit-identical).
The only communication is ONE gather of the final-level bands to rank 0 (NCCL over NVLink on the
GPU box, gloo in the CPU tests) — the reference itself is single-GPU (src/main.rs:450-460).
"""
from __future__ import annotations

import torch
import torch.distributed as dist

BAND_ALIGN = 8   # bands are whole 8-row workgroup rows of the reference dispatch (compute.glsl:5)


def band_rows(height: int, world: int, rank: int, align: int = BAND_ALIGN):
    """Rows [y0, y1) of the final level owned by `rank`: multiples of `align`, sizes differ by <= align."""
    units = (height + align - 1) // align
    base, extra = divmod(units, world)
    u0 = rank * base + min(rank, extra)
    u1 = u0 + base + (1 if rank < extra else 0)
    return min(u0 * align, height), min(u1 * align, height)


def all_bands(height: int, world: int, align: int = BAND_ALIGN):
    return [band_rows(height, world, r, align) for r in range(world)]


def gather_bands(local: torch.Tensor, height: int, dst: int = 0, group=None):
    """Gather per-rank bands `local` = [..., rows_r, width] into [..., height, width] on rank `dst`.

    One collective: every rank contributes a band padded to the largest band height
    (dist.gather needs equal shapes); rank `dst` strips the padding.  Returns the full image on
    `dst`, None elsewhere.
    """
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    bands = all_bands(height, world)
    max_rows = max(y1 - y0 for (y0, y1) in bands)
    y0, y1 = bands[rank]
    assert local.shape[-2] == y1 - y0, (local.shape, bands[rank])
    lead, width = local.shape[:-2], local.shape[-1]
    if y1 - y0 == max_rows:
        send = local.contiguous()
    else:
        send = local.new_zeros(*lead, max_rows, width)
        send[..., : y1 - y0, :] = local
    if rank == dst:
        parts = [torch.empty_like(send) for _ in range(world)]
        dist.gather(send, parts, dst=dst, group=group)
        full = local.new_empty(*lead, height, width)
        for (b0, b1), p in zip(bands, parts):
            full[..., b0:b1, :] = p[..., : b1 - b0, :]
        return full
    dist.gather(send, None, dst=dst, group=group)
    return None
