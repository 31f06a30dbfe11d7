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


def weighted_bands(row_weights, world: int, align: int = BAND_ALIGN):
    """Contiguous row bands of (nearly) equal total weight: [(y0, y1)] * world, boundaries on multiples of `align`,
    every band at least `align` rows.  `row_weights[y]` is the estimated cost of final-level row y (e.g.
    `row_cost` below).  Pure function of its arguments, so every rank computes the same partition from the same
    pilot frame without talking to the others.  Equal weights reproduce `all_bands` up to rounding."""
    import numpy as np
    w = np.asarray(row_weights, dtype=np.float64).reshape(-1)
    height = int(w.shape[0])
    units = (height + align - 1) // align
    if units < world:
        raise ValueError(f"{height} rows cannot be split into {world} bands of {align}-row units")
    pad = np.zeros(units * align, dtype=np.float64)
    pad[:height] = np.maximum(w, 0.0)
    uw = pad.reshape(units, align).sum(axis=1) + 1e-12   # strictly increasing prefix sums
    cum = np.concatenate([[0.0], np.cumsum(uw)])
    cuts = [0]
    for k in range(1, world):
        target = cum[-1] * k / world
        u = int(np.searchsorted(cum, target))            # first prefix >= target
        if u > 0 and target - cum[u - 1] < cum[u] - target:
            u -= 1                                       # nearer boundary
        u = max(u, cuts[-1] + 1)                         # at least one unit per band ...
        u = min(u, units - (world - k))                  # ... for the bands still to come, too
        cuts.append(u)
    cuts.append(units)
    return [(min(a * align, height), min(b * align, height)) for a, b in zip(cuts, cuts[1:])]


def rebalanced_bands(bands, times, height: int, align: int = BAND_ALIGN):
    """One step of measured-cost balancing: given the bands of a trial and the time each rank took for its band,
    cut new bands of equal PREDICTED time, taking the cost of a row to be uniform inside its old band
    (t_r / rows_r).  Every rank holds the same `times` (gathered), so every rank computes the same cuts.  The
    estimate ignores the part of a rank's time that does not scale with its rows (the coarse levels), so the caller
    times the result and keeps it only if it is faster."""
    import numpy as np
    assert len(bands) == len(times) and bands[0][0] == 0 and bands[-1][1] == height
    w = np.zeros(height, dtype=np.float64)
    for (y0, y1), t in zip(bands, times):
        if y1 > y0:
            w[y0:y1] = max(float(t), 0.0) / (y1 - y0)
    return weighted_bands(w, len(bands), align)


def row_cost(final_level: torch.Tensor, render_dist: float, hit_weight: float):
    """Estimated cost of each final-level row from one traced frame [rows, width]: a texel that reached
    RENDER_DIST is background (resolved by the far-miss shortcut: a load, a compare and a store), anything else
    is marched (second trips, tails); `hit_weight` = cost of a marched texel in units of a background texel."""
    marched = (final_level < render_dist).sum(dim=-1).to(torch.float64)
    return (final_level.shape[-1] + (hit_weight - 1.0) * marched).cpu().numpy()


def gather_bands(local: torch.Tensor, height: int, dst: int = 0, group=None, bands=None, out=None):
    """Gather per-rank bands `local` = [..., rows_r, width] into [..., height, width] on rank `dst`.

    `bands` = [(y0, y1)] per rank (default: the equal split `all_bands`).  One exchange step: a row band of a
    row-major image is contiguous, so when there is a single leading image every rank sends exactly its rows and
    rank `dst` receives them in place (batched point-to-point, no padding, no copy).  With several leading images the
    bands are padded to the largest band height and gathered with one dist.gather.  `out` (rank `dst`, single image):
    a preallocated [..., height, width] tensor to receive into.  Returns the full image on `dst`, None elsewhere.
    """
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if bands is None:
        bands = all_bands(height, world)
    assert len(bands) == world and bands[0][0] == 0 and bands[-1][1] == height, bands
    y0, y1 = bands[rank]
    assert local.shape[-2] == y1 - y0, (local.shape, bands[rank])
    lead, width = local.shape[:-2], local.shape[-1]
    n_lead = 1
    for d in lead:
        n_lead *= int(d)
    if n_lead == 1:
        if rank == dst:
            full = out if out is not None else local.new_empty(*lead, height, width)
            flat = full.view(height, width)
            flat[y0:y1] = local.reshape(y1 - y0, width)
            ops = [dist.P2POp(dist.irecv, flat[b0:b1], r, group) for r, (b0, b1) in enumerate(bands) if r != dst and b1 > b0]
            if ops:
                for req in dist.batch_isend_irecv(ops):
                    req.wait()
            return full
        if y1 > y0:
            for req in dist.batch_isend_irecv([dist.P2POp(dist.isend, local.reshape(y1 - y0, width).contiguous(), dst, group)]):
                req.wait()
        return None
    max_rows = max(b1 - b0 for (b0, b1) in bands)
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
