"""Multi-GPU host logic on CPU: the row-block partition and the readout reduce, world_size 2
over gloo.  The device side of the same partition is tested on the GPU (test_shards_sum_to_single)."""
import os
import socket
import sys

import numpy as np
import pytest

import common
from elevenrender_b200 import sharding


@pytest.mark.parametrize("count,rows,y_res", [(1, 8, 1080), (2, 8, 1080), (3, 4, 50), (8, 8, 1080), (8, 4, 7), (5, 12, 2160)])
def test_partition_covers_every_row_once(count, rows, y_res):
    assert sharding.check_partition(count, y_res, rows)
    px = sum(sharding.owned_pixel_count(k, count, 1920, y_res, rows) for k in range(count))
    assert px == 1920 * y_res
    for y in (0, rows - 1, rows, y_res - 1):
        k = sharding.owner_of_row(y, count, rows)
        assert sharding.owned_rows(k, count, y_res, rows)[y]


def test_work_tickets_match_kernel_mapping():
    """tickets = 8x4 tiles over the owned row blocks, column-major inside a block, x padded to 8
    (er_ticket_pixel): replay the mapping"""
    for (count, rows, x_res, y_res) in ((1, 8, 48, 32), (2, 8, 1920, 1080), (3, 4, 50, 30), (8, 8, 20, 100)):
        seen = np.zeros((y_res, x_res), dtype=np.int32)
        for k in range(count):
            n = sharding.work_tickets(k, count, x_res, y_res, rows)
            w = np.arange(n, dtype=np.int64)
            tiles_x = -(-x_res // 8)
            per_block = tiles_x * (rows // 4) * 32
            ob, r = w // per_block, w % per_block
            tile, lane = r >> 5, r & 31
            tiles_y = rows // 4
            tx, ty = tile // tiles_y, tile % tiles_y  # column by column inside a row block
            x = tx * 8 + (lane & 7)
            y = (ob * count + k) * rows + ty * 4 + (lane >> 3)
            ok = (x < x_res) & (y < y_res)
            np.add.at(seen, (y[ok], x[ok]), 1)
            assert sharding.owned_rows(k, count, y_res, rows)[y[ok]].all()
        assert (seen == 1).all()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, golden_path, rows, out_dir):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = np.load(golden_path)
        full = g["passes_4"]  # (4 passes, h, w, 4) produced by the compiled reference
        # what this shard's device buffers hold: its rows, exact zeros elsewhere
        mine = np.stack([sharding.mask_to_shard(p, rank, world, rows) for p in full])
        buf = torch.from_numpy(mine.copy())
        sharding.reduce_pass(buf, dst=0)
        n = sharding.min_samples(5 if rank == 0 else 7)
        tmax = sharding.max_over_ranks([1.0 + rank, 10.0 - rank])
        if rank == 0:
            np.savez(os.path.join(out_dir, "reduced.npz"), frame=buf.numpy(), n=n, tmax=np.array(tmax))
        else:
            assert n == 5 and tmax == [float(world), 10.0]
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("rows", [8, 4])
def test_readout_reduce_gloo_world2(tmp_path, rows):
    """two ranks, each holding its row blocks of the reference's frame: the reduce on rank 0 is the
    frame, bit for bit; sample count = min over shards; step time = max over shards"""
    import torch.multiprocessing as mp
    golden = os.path.join(common.GOLDEN_DIR, "textured.npz")
    port = _free_port()
    mp.spawn(_worker, args=(2, port, golden, rows, str(tmp_path)), nprocs=2, join=True)
    r = np.load(tmp_path / "reduced.npz")
    want = np.load(golden)["passes_4"]
    assert np.array_equal(r["frame"].view(np.uint32), want.view(np.uint32))
    assert int(r["n"]) == 5
    assert list(r["tmax"]) == [2.0, 10.0]


def test_single_process_is_identity():
    import torch
    buf = torch.arange(12, dtype=torch.float32)
    assert sharding.reduce_pass(buf.clone()).equal(buf)
    assert sharding.min_samples(9) == 9
    assert sharding.max_over_ranks([1.5, 2.5]) == [1.5, 2.5]


# ---- strong scaling: cost-balanced contiguous stripes + owned-rows gather into a shared frame -----------
def test_balance_stripes_splits_cost_evenly():
    from elevenrender_b200 import capi
    rng = np.random.default_rng(3)
    for (y_res, rows, n) in ((1080, 8, 8), (1080, 8, 2), (272, 8, 4), (50, 4, 3), (2160, 8, 8)):
        # a sky band (cheap), a dense band (10x), sampled on every 4th row block only
        cost = np.zeros(y_res, dtype=np.uint32)
        dens = np.where(np.arange(y_res) < y_res // 3, 100, 1000) + rng.integers(0, 50, y_res)
        nb = -(-y_res // rows)
        for b in range(0, nb, 4):
            cost[b * rows:(b + 1) * rows] = dens[b * rows:(b + 1) * rows]
        yb = capi.balance_stripes(cost, rows, n)
        assert yb[0] == 0 and yb[-1] == y_res and np.all(np.diff(yb) >= 0)
        assert all(int(v) % rows == 0 for v in yb[:-1])
        per = np.array([dens[yb[k]:yb[k + 1]].sum() for k in range(n)], dtype=np.float64)
        # every stripe within one row block's cost of the mean
        assert np.abs(per - per.mean()).max() <= 1050 * rows * 1.5, (yb, per)
        masks = sum(sharding.stripe_rows(list(yb), k, y_res).astype(int) for k in range(n))
        assert (masks == 1).all()
    # degenerate inputs: more shards than row blocks, all-zero cost
    yb = capi.balance_stripes(np.zeros(16, dtype=np.uint32), 8, 4)
    assert yb[0] == 0 and yb[-1] == 16 and np.all(np.diff(yb) >= 0)


def _stripe_worker(rank, world, port, golden_path, out_dir):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        full = np.load(golden_path)["passes_4"][0]  # the reference's beauty frame (h, w, 4)
        h, w = full.shape[:2]
        # each rank sampled different row blocks in its cost prepass; the sum over ranks is the cost profile
        cost = np.zeros(h, dtype=np.uint32)
        for b in range(rank, -(-h // 4), world):
            cost[b * 4:(b + 1) * 4] = 10 + 5 * b
        yb = sharding.balanced_stripes(cost, world, 4)
        frame = sharding.SharedFrame(f"test_{port}", h, w, rank, dist.barrier, pin=False)
        if rank == 0:
            frame.array[...] = -1.0
        dist.barrier()
        own = sharding.stripe_rows(yb, rank, h)
        frame.array[own] = full[own]  # what er_get_pass_owned does with this rank's device rows
        dist.barrier()
        if rank == 0:
            np.savez(os.path.join(out_dir, "gathered.npz"), frame=frame.array.copy(), yb=np.array(yb))
        frame.close()
    finally:
        dist.destroy_process_group()


def test_stripe_gather_gloo_world2(tmp_path):
    """two ranks agree on cost-balanced stripes (all-reduce of their prepass shares) and each writes its rows
    into the shared frame: rank 0 then holds the reference's frame bit for bit"""
    import torch.multiprocessing as mp
    golden = os.path.join(common.GOLDEN_DIR, "textured.npz")
    port = _free_port()
    mp.spawn(_stripe_worker, args=(2, port, golden, str(tmp_path)), nprocs=2, join=True)
    r = np.load(tmp_path / "gathered.npz")
    want = np.load(golden)["passes_4"][0]
    assert np.array_equal(r["frame"].view(np.uint32), want.view(np.uint32))
    yb = r["yb"]
    assert yb[0] == 0 and yb[-1] == want.shape[0] and 0 < yb[1] < want.shape[0]


def test_rebalance_stripes_feedback_converges():
    """er_rebalance_stripes: boundaries stay block-aligned, monotonic, every stripe keeps a block, and iterating on
    a time model with a fixed part per launch (the longest pixels of a stripe) settles at equal times"""
    from elevenrender_b200 import capi
    y_res, rows, n = 1080, 8, 8
    dens = np.where(np.arange(y_res) < 300, 1.0, 3.0) + 0.5 * np.sin(np.arange(y_res) / 40.0)  # ms per row

    def times(yb):
        return np.array([100.0 + dens[yb[k]:yb[k + 1]].sum() for k in range(n)])

    yb = np.array([0, 136, 272, 408, 544, 680, 816, 952, 1080], dtype=np.int32)
    first = times(yb)
    for _ in range(4):
        yb = capi.rebalance_stripes(yb, times(yb), rows)
        assert yb[0] == 0 and yb[-1] == y_res and np.all(np.diff(yb) >= rows)
        assert all(int(v) % rows == 0 for v in yb[:-1])
    last = times(yb)
    assert last.max() < first.max() and last.max() - last.min() <= 2 * 3.5 * rows, (yb, last)
    # equal times: nothing moves; zero times: unchanged; one stripe with all the time: the others keep one block
    same = capi.rebalance_stripes([0, 536, 1080], [5.0, 5.0], 8)
    assert list(same) == [0, 536, 1080]
    assert list(capi.rebalance_stripes([0, 536, 1080], [0.0, 0.0], 8)) == [0, 536, 1080]
    moved = capi.rebalance_stripes([0, 8, 16, 1080], [0.0, 0.0, 9.0], 8)  # proportional: 368, 728; over-relaxed 1.5 x
    assert moved[0] == 0 and moved[3] == 1080 and moved[1] == 552 and moved[2] == 1072
    squeezed = capi.rebalance_stripes([0, 1064, 1072, 1080], [1000.0, 0.0, 0.0], 8)
    assert np.all(np.diff(squeezed) >= 8) and squeezed[-1] == 1080
    with pytest.raises(capi.ErError):
        capi.rebalance_stripes([0, 536, 1080], [1.0, float("nan")], 8)


def _rebalance_worker(rank, world, port, out_dir):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        yb = sharding.rebalanced_stripes([0, 136, 272], 300.0 if rank == 0 else 120.0, rank, 8)
        np.save(os.path.join(out_dir, f"yb{rank}.npy"), np.array(yb))
    finally:
        dist.destroy_process_group()


def test_rebalanced_stripes_gloo_world2(tmp_path):
    """two ranks contribute the time of their own stripe and arrive at the same new boundaries"""
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_rebalance_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    a, b = np.load(tmp_path / "yb0.npy"), np.load(tmp_path / "yb1.npy")
    assert np.array_equal(a, b) and a[0] == 0 and a[2] == 272
    # 210 of the 420 ms lie 0.7 into the first stripe's 136 rows (95.2); moved 1.5 x as far from 136: 74.8 -> block edge 72
    assert a[1] == 72


def test_order_blocks_by_cost():
    """er_order_blocks: descending cost, unsampled blocks take the nearest sampled block's cost, stable for ties"""
    from elevenrender_b200 import capi
    c = np.zeros(64, dtype=np.uint32)
    c[0:8], c[32:40], c[56:64] = 5, 50, 20
    assert list(capi.order_blocks(c, 8, 0, 64)) == [3, 4, 5, 6, 7, 0, 1, 2]
    assert list(capi.order_blocks(c, 8, 16, 64)) == [1, 2, 3, 4, 5, 0]
    assert list(capi.order_blocks(c, 8, 16, 60)) == [1, 2, 3, 4, 5, 0]  # a partial last block counts
    assert list(capi.order_blocks(np.zeros(32, dtype=np.uint32), 8, 0, 32)) == [0, 1, 2, 3]
    assert len(capi.order_blocks(c, 8, 8, 8)) == 0
    with pytest.raises(capi.ErError):
        capi.order_blocks(c, 8, 4, 64)  # not at a block edge
