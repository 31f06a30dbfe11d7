"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: row-band partition + the single gather.
Each rank traces ITS band with the CPU checker (standing in for its GPU) through the same
sharding.gather_bands the bench uses; rank 0 must hold an image bit-identical to the 1-rank image."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from raytracing_engine_b200 import sharding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_band_rows_cover_image_in_multiples_of_8():
    for H in (600, 1080, 2160, 4320, 48, 8, 136):
        for world in (1, 2, 3, 4, 8):
            bands = sharding.all_bands(H, world)
            assert bands[0][0] == 0 and bands[-1][1] == H
            for (a0, a1), (b0, b1) in zip(bands, bands[1:]):
                assert a1 == b0
            for (y0, y1) in bands:
                assert y0 % 8 == 0 and (y1 % 8 == 0 or y1 == H)
            sizes = [y1 - y0 for y0, y1 in bands]
            assert max(sizes) - min(sizes) <= 8
    assert [y1 - y0 for y0, y1 in sharding.all_bands(4320, 8)] == [544, 544, 544, 544, 536, 536, 536, 536]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, W, H, frames, q):
    sys.path.insert(0, ROOT)
    import oracle
    from raytracing_engine_b200 import camera, scene
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        o = oracle.load("port")
        sph = scene.spheres_of(scene.default_mutable_data())
        n = o.level_count(float(W))
        Hf = o.level_dims(n, float(W), float(H))[-1][1]
        Wf = o.level_dims(n, float(W), float(H))[-1][0]
        y0, y1 = sharding.band_rows(Hf, world, rank)
        rots, poss = camera.bench_camera_path(frames)
        local = []
        for k in range(frames):
            lv, _ = o.trace(sph, [W, H], scene.view_ratio(W, H), rots[k], poss[k], 1000.0, n,
                            rect=[0, y0, Wf, y1], threads=1)
            local.append(torch.from_numpy(lv[-1][y0:y1].copy()))
        full = sharding.gather_bands(torch.stack(local), Hf, dst=0)
        if rank == 0:
            ok = True
            for k in range(frames):
                ref, _ = o.trace(sph, [W, H], scene.view_ratio(W, H), rots[k], poss[k], 1000.0, n, threads=1)
                ok = ok and np.array_equal(full[k].numpy().view(np.uint32), ref[-1].view(np.uint32))
            q.put(bool(ok))
        else:
            assert full is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,W,H", [(2, 320, 200), (3, 250, 130)])
def test_banded_trace_plus_gather_is_bit_identical(world, W, H):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, W, H, 2, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


def test_rebalanced_bands_move_rows_to_the_faster_ranks():
    """Measured-cost balancing (bench.py strong_c4): ranks that took longer give rows away, cuts stay on multiples of 8,
    the bands still tile the image, and equal times change nothing beyond rounding."""
    H = 4320
    bands = sharding.all_bands(H, 8)
    times = [26.3, 26.0, 26.3, 26.2, 25.4, 25.2, 24.6, 23.6]          # the N = 8 run of round 2
    nb = sharding.rebalanced_bands(bands, times, H)
    assert nb[0][0] == 0 and nb[-1][1] == H and all(a[1] == b[0] for a, b in zip(nb, nb[1:]))
    assert all(y0 % 8 == 0 and y1 > y0 for y0, y1 in nb)
    assert nb[0][1] - nb[0][0] < bands[0][1] - bands[0][0] and nb[-1][1] - nb[-1][0] > bands[-1][1] - bands[-1][0]
    pred = [t / (b[1] - b[0]) * (n[1] - n[0]) for t, b, n in zip(times, bands, nb)]   # crude: same per-row cost per rank
    assert max(pred) < max(times)
    same = sharding.rebalanced_bands(bands, [5.0] * 8, H)
    assert all(abs(a[0] - b[0]) <= 8 and abs(a[1] - b[1]) <= 8 for a, b in zip(same, bands))


def test_weighted_bands_balance_cost_in_multiples_of_8():
    rs = np.random.RandomState(7)
    for H in (136, 600, 1080, 4320):
        for world in (2, 3, 8):
            w = np.ones(H)
            a = int(rs.randint(0, H // 2))
            w[a:a + H // 4] += rs.uniform(2, 30)
            bands = sharding.weighted_bands(w, world)
            assert len(bands) == world and bands[0][0] == 0 and bands[-1][1] == H
            for (a0, a1), (b0, b1) in zip(bands, bands[1:]):
                assert a1 == b0
            for (y0, y1) in bands:
                assert y1 > y0 and y0 % 8 == 0 and (y1 % 8 == 0 or y1 == H)
            cost = [w[y0:y1].sum() for y0, y1 in bands]
            # no band is heavier than the ideal share by more than one 8-row unit of the heaviest rows
            assert max(cost) <= w.sum() / world + 8 * w.max() + 1e-9
    assert sharding.weighted_bands(np.ones(4320), 8)[3] == (1616, 2160)
    with pytest.raises(ValueError):
        sharding.weighted_bands(np.ones(16), 3)


def _worker_weighted(rank, world, port, W, H, q):
    """Cost-balanced bands cut from a pilot frame + the in-place point-to-point gather of ONE accumulated image."""
    sys.path.insert(0, ROOT)
    import oracle
    from raytracing_engine_b200 import camera, scene
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        o = oracle.load("port")
        sph = scene.spheres_of(scene.default_mutable_data())
        n = o.level_count(float(W))
        Wf, Hf = o.level_dims(n, float(W), float(H))[-1]
        rots, poss = camera.static_path(1)
        args = (sph, [W, H], scene.view_ratio(W, H), rots[0], poss[0], 1000.0, n)
        pilot, _ = o.trace(*args, threads=1)
        bands = sharding.weighted_bands(sharding.row_cost(torch.from_numpy(pilot[-1]), 1000.0, 8.0), world)
        y0, y1 = bands[rank]
        lv, _ = o.trace(*args, rect=[0, y0, Wf, y1], threads=1)
        local = torch.from_numpy(lv[-1][y0:y1].copy())[None]
        full = sharding.gather_bands(local, Hf, dst=0, bands=bands)
        if rank == 0:
            sizes = [b1 - b0 for b0, b1 in bands]
            q.put(bool(np.array_equal(full[0].numpy().view(np.uint32), pilot[-1].view(np.uint32))) and max(sizes) > min(sizes))
        else:
            assert full is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_weighted_bands_plus_in_place_gather_is_bit_identical(world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_weighted, args=(r, world, port, 320, 200, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True
