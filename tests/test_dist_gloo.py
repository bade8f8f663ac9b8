"""
Host-side logic of the multi-GPU path (fv-srn_b200/dist.py) on CPU: world_size-2 (and 3) gloo process groups, a
deterministic stand-in renderer. The GPU path uses the same functions with NCCL (bench.py --gpus N).
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import fvsrn_b200  # noqa: F401
from fvsrn_b200 import dist as fdist


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _fake_frame(width, height, batch=1):
    """Deterministic (B,8,H,W) 'render' every rank can compute: value = f(b, c, y, x)."""
    b, c, y, x = np.meshgrid(np.arange(batch), np.arange(8), np.arange(height), np.arange(width), indexing="ij")
    return torch.from_numpy((b * 1000 + c * 100 + y + x / 4096.0).astype(np.float32))


def _worker(rank, world, port, width, height, band, views, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        full = _fake_frame(width, height, 2)

        def render_rect(rect):                       # a renderer that only fills the requested rectangle
            x0, y0, w, h = rect
            img = torch.zeros_like(full)
            img[:, :, y0:y0 + h, x0:x0 + w] = full[:, :, y0:y0 + h, x0:x0 + w]
            return img

        frame = fdist.render_frame_sharded(render_rect, width, height, band)
        ok_frame = (frame is None) if rank != 0 else bool(torch.equal(frame, full[:, :4]))

        def render_view(i):
            return _fake_frame(16, 8)[:, :4] + float(i)

        sweep = fdist.render_views_sharded(render_view, views)
        if rank == 0:
            want = torch.cat([render_view(i) for i in range(views)], dim=0)
            ok_views = bool(torch.equal(sweep, want))
        else:
            ok_views = sweep is None
        if views % world == 0:      # even split: pieces land directly in a caller-provided buffer, both collectives
            per = views // world
            mine = torch.cat([render_view(i) for i in range(rank * per, (rank + 1) * per)], dim=0)
            for coll in ("gather", "all_gather"):
                buf = torch.full((views,) + tuple(mine.shape[1:]), -1.0)
                got = fdist.gather_views(mine, views, out=buf, collective=coll)
                if rank == 0:
                    ok_views = ok_views and got is buf and bool(torch.equal(buf, want))
                else:
                    ok_views = ok_views and got is None
        # cost-balanced contiguous row bands of ONE packed-pixel frame (RGBA8 / half4), received in place on rank 0, twice
        # with double-buffered asynchronous transfers (the overlap bench.py --sharding tiles uses)
        cost = np.linspace(0.0, 3.0, height) ** 2
        bands = fdist.partition_rows_by_cost(cost, world, align=8)
        for dtype, tail in ((torch.int32, ()), (torch.float16, (4,))):
            frames = [(torch.arange(height * width * (4 if tail else 1)).reshape((1, height, width) + tail) % 251 + k).to(dtype) for k in range(3)]
            bufs = [torch.full_like(frames[0], -1) for _ in range(2)] if rank == 0 else [None, None]
            pending = [None, None]
            for k, f in enumerate(frames):
                slot = k % 2
                if pending[slot] is not None:
                    for w in pending[slot][0]: w.wait()
                    if rank == 0: ok_frame = ok_frame and bool(torch.equal(bufs[slot], pending[slot][1]))
                y0, h = bands[rank]
                local = f[:, y0:y0 + h].clone()
                if rank == 0: bufs[slot][:, y0:y0 + h] = local
                pending[slot] = (fdist.gather_row_bands(local, bands, bufs[slot], async_op=True), f)
            for slot in range(2):
                if pending[slot] is not None:
                    for w in pending[slot][0]: w.wait()
                    if rank == 0: ok_frame = ok_frame and bool(torch.equal(bufs[slot], pending[slot][1]))
        if views % world == 0:      # asynchronous view gather: (result, work handle)
            per = views // world
            mine = torch.cat([render_view(i) for i in range(rank * per, (rank + 1) * per)], dim=0)
            buf = torch.full((views,) + tuple(mine.shape[1:]), -1.0)
            got, work = fdist.gather_views(mine, views, out=buf, async_op=True)
            work.wait()
            ok_views = ok_views and ((got is buf and bool(torch.equal(buf, want))) if rank == 0 else got is None)
        q.put((rank, ok_frame, ok_views))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,width,height,band,views", [(2, 64, 200, 64, 5), (3, 40, 136, 8, 7), (2, 32, 24, 64, 2), (2, 32, 24, 8, 6)])
def test_sharded_frame_and_sweep_gloo(world, width, height, band, views):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, width, height, band, views, q)) for r in range(world)]
    for p in procs: p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok_f and ok_v for _, ok_f, ok_v in res), res


def test_partitions_cover_everything_exactly_once():
    for world in (1, 2, 3, 4, 8):
        for n in (1, 7, 8, 64):
            got = sum((fdist.partition_views(n, world, r) for r in range(world)), [])
            assert got == list(range(n))
            sizes = [len(fdist.partition_views(n, world, r)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
        for (w, h, band) in ((2048, 2048, 64), (100, 90, 8), (64, 8, 64)):
            cover = np.zeros((h, w), np.int32)
            for r in range(world):
                for (x0, y0, ww, hh) in fdist.partition_bands(w, h, world, r, band):
                    assert x0 == 0 and ww == w and y0 % 8 == 0
                    cover[y0:y0 + hh, x0:x0 + ww] += 1
            assert (cover == 1).all()
        for h in (8, 90, 2048):
            cost = np.abs(np.sin(np.arange(h) / 37.0)) * (np.arange(h) > h // 5)
            bands = fdist.partition_rows_by_cost(cost, world, align=8)
            assert len(bands) == world and bands[0][0] == 0 and sum(b[1] for b in bands) == h
            assert all(bands[i][0] + bands[i][1] == bands[i + 1][0] for i in range(world - 1))
            assert all(y0 % 8 == 0 for y0, _ in bands)
            if h == 2048 and world > 1:     # equal cost within the rounding to 8 rows
                shares = [cost[y0:y0 + hh].sum() / cost.sum() for y0, hh in bands]
                assert max(shares) - min(shares) < 0.02
    with pytest.raises(ValueError):
        fdist.partition_bands(64, 64, 2, 0, band=12)
    with pytest.raises(ValueError):
        fdist.partition_views(4, 2, 2)


def test_rebalance_rows_by_time_converges_on_the_true_cost():
    """Feedback rebalancing of the row bands: the estimate sees the box (a smooth bump), the true cost has structure it
    does not know (here: rays of the lower half terminate early at 40 % of their length). Two measured iterations bring
    every band within a few percent of the mean."""
    H, world = 2048, 8
    y = np.arange(H)
    estimate = np.clip(1.0 - ((y - H / 2) / (H / 2)) ** 2, 0, None) + 0.05
    true = estimate * np.where(y > H // 2, 0.4, 1.0)
    bands = fdist.partition_rows_by_cost(estimate, world, align=8)
    measure = lambda bs: [float(true[y0:y0 + h].sum()) for y0, h in bs]
    t0 = measure(bands)
    assert max(t0) / np.mean(t0) > 1.3                      # the estimate alone is badly balanced on this frame
    cost = estimate
    for _ in range(2):
        bands, cost = fdist.rebalance_rows_by_time(cost, bands, measure(bands), align=8)
        assert bands[0][0] == 0 and sum(h for _, h in bands) == H and all(y0 % 8 == 0 for y0, _ in bands)
        assert all(bands[i][0] + bands[i][1] == bands[i + 1][0] for i in range(world - 1))
    t2 = measure(bands)
    assert max(t2) / np.mean(t2) < 1.06                  # what is left is the rounding of the cuts to 8 rows (of ~130 per band here)
    # an idle rank (empty band) and zero measurements do not break it
    b2, _ = fdist.rebalance_rows_by_time(estimate, [(0, H)] + [(H, 0)] * 3, [5.0, 0.0, 0.0, 0.0], align=8)
    assert sum(h for _, h in b2) == H and all(h > 0 for _, h in b2)


def test_single_process_paths_need_no_process_group():
    full = _fake_frame(48, 40)
    rects = fdist.partition_bands(48, 40, 1, 0, 16)
    packed = fdist.pack_bands(full, rects)
    assert packed.shape == (1, 4, 40, 48)
    assert torch.equal(fdist.gather_bands(packed, 48, 40, 16), full[:, :4])
    assert fdist.world_info() == (0, 1)
