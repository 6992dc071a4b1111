"""Multi-GPU host logic: pixel row-block shards and the readout collective.

One process per GPU.  Shard k of G renders the row blocks  b  with  b % G == k  (block = `rows_per_block`
image rows, the partition er_set_shard installs on the device side: er_render.cuh er_owns_row /
er_ticket_pixel); the scene is replicated.  Non-owned pixels of a shard's pass buffer hold exact zeros
(alpha included), so the frame is the element-wise SUM over shards — x + 0 is exact in fp32, hence the
G-GPU frame is bit-identical to the 1-GPU frame.  There is no collective inside the render loop; the only
exchange is this reduce at readout (RenderingManager::get_pass, src/Managers.cpp:287-302) and the
min over shards of the completed sample count (get_render_info, :211-232).

Strong scaling (one frame over N GPUs) uses CONTIGUOUS STRIPES instead: shard k owns rows
[y[k], y[k+1]) with the boundaries placed so that the stripes cost the same (er_row_cost prepass, summed
over the ranks, er_balance_stripes); and the readout is a gather of owned rows, not a reduce of zeros:
every rank's GPU writes its rows straight into ONE page-locked frame in POSIX shared memory
(er_get_pass_owned into SharedFrame), N PCIe links in parallel.

Everything here works on CPU tensors with the gloo backend (tests/test_sharding_cpu.py) and on CUDA
tensors with NCCL (bench.py, the N>1 readout).
"""
import mmap
import os

import numpy as np

ROWS_PER_BLOCK = 8


def owner_of_row(y, shard_count, rows_per_block=ROWS_PER_BLOCK):
    """shard that owns image row y (same formula as er_owns_row)"""
    return (y // rows_per_block) % shard_count


def owned_rows(shard_index, shard_count, y_res, rows_per_block=ROWS_PER_BLOCK):
    """boolean mask over the y_res rows of the frame"""
    y = np.arange(y_res)
    return ((y // rows_per_block) % shard_count) == shard_index


def owned_pixel_count(shard_index, shard_count, x_res, y_res, rows_per_block=ROWS_PER_BLOCK):
    return int(owned_rows(shard_index, shard_count, y_res, rows_per_block).sum()) * x_res


def work_tickets(shard_index, shard_count, x_res, y_res, rows_per_block=ROWS_PER_BLOCK):
    """number of pixel tickets the render kernel hands out for this shard: 8x4-pixel tiles over the
    shard's row blocks, padded in x to a multiple of 8 (er_ticket_pixel); rows_per_block % 4 == 0"""
    assert rows_per_block % 4 == 0
    blocks = -(-y_res // rows_per_block)
    own_blocks = len(range(shard_index, blocks, shard_count))
    tiles_x = -(-x_res // 8)
    return own_blocks * tiles_x * (rows_per_block // 4) * 32


def check_partition(shard_count, y_res, rows_per_block=ROWS_PER_BLOCK):
    """every row has exactly one owner"""
    total = np.zeros(y_res, dtype=np.int64)
    for k in range(shard_count):
        total += owned_rows(k, shard_count, y_res, rows_per_block)
    return bool((total == 1).all())


def mask_to_shard(frame, shard_index, shard_count, rows_per_block=ROWS_PER_BLOCK):
    """what a shard's pass buffer holds given the full (h, w, 4) frame: its rows, exact zeros elsewhere"""
    own = owned_rows(shard_index, shard_count, frame.shape[0], rows_per_block)
    out = np.zeros_like(frame)
    out[own] = frame[own]
    return out


def reduce_pass(buf, dst=0, group=None):
    """sum of the shards' pass buffers onto rank `dst` (in place on dst).  `buf` is a torch tensor
    (CUDA -> NCCL over NVLink, CPU -> gloo).  Exact: each pixel is non-zero on one shard only."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(buf, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return buf


def min_samples(samples, group=None):
    """get_info across shards: a frame has n samples when every shard has n"""
    import torch
    import torch.distributed as dist
    if not (dist.is_initialized() and dist.get_world_size(group) > 1):
        return int(samples)
    t = torch.tensor([int(samples)], dtype=torch.int64)
    if dist.get_backend(group) == "nccl":
        t = t.cuda()
    dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
    return int(t.item())


def max_over_ranks(values, group=None):
    """device-timed durations: the job takes as long as its slowest shard"""
    import torch
    import torch.distributed as dist
    t = torch.tensor([float(v) for v in values], dtype=torch.float64)
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        if dist.get_backend(group) == "nccl":
            t = t.cuda()
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return [float(v) for v in t.cpu()]


# ---- contiguous cost-balanced stripes + shared-memory frame gather (strong scaling) -----------------
def balanced_stripes(cost_per_row, shard_count, rows_per_block=ROWS_PER_BLOCK, group=None, return_cost=False):
    """stripe boundaries y[0..shard_count] from this rank's share of the cost prepass: the per-row costs are
    summed over the ranks (each rank sampled different row blocks), then split evenly by er_balance_stripes
    (host-only C function: the C++ host and this mirror share it)"""
    import torch
    import torch.distributed as dist
    from . import capi
    t = torch.from_numpy(np.ascontiguousarray(cost_per_row).astype(np.int64))
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        if dist.get_backend(group) == "nccl":
            t = t.cuda()
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    cost = np.minimum(t.cpu().numpy(), 0xFFFFFFFF).astype(np.uint32)
    yb = [int(v) for v in capi.balance_stripes(cost, rows_per_block, shard_count)]
    return (yb, cost) if return_cost else yb


def rebalanced_stripes(bounds, my_ms, shard_index, rows_per_block=ROWS_PER_BLOCK, group=None):
    """feedback step of the balance (one process per GPU): every rank contributes the kernel time its stripe took,
    all ranks get the same new boundaries (er_rebalance_stripes)"""
    import torch
    import torch.distributed as dist
    from . import capi
    n = len(bounds) - 1
    t = torch.zeros(n, dtype=torch.float64)
    t[shard_index] = float(my_ms)
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        if dist.get_backend(group) == "nccl":
            t = t.cuda()
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return [int(v) for v in capi.rebalance_stripes(bounds, t.cpu().numpy(), rows_per_block)]


def stripe_rows(bounds, shard_index, y_res):
    """boolean mask over the rows of the frame: the rows of stripe shard_index"""
    y = np.arange(y_res)
    return (y >= bounds[shard_index]) & (y < bounds[shard_index + 1])


class SharedFrame:
    """One (h, w, 4) float32 frame in POSIX shared memory, mapped by every rank of the node (and page-locked
    in place when `pin`), so that each rank's GPU writes the rows it owns straight into the frame rank 0 reads.
    `barrier` is a callable that synchronises the ranks (dist.barrier)."""

    def __init__(self, tag, h, w, rank, barrier, pin=True):
        self.path = f"/dev/shm/er_frame_{tag}_{h}x{w}"
        self.nbytes = h * w * 16
        if rank == 0:
            with open(self.path, "wb") as f:
                f.truncate(self.nbytes)
        barrier()
        self.f = open(self.path, "r+b")
        self.mm = mmap.mmap(self.f.fileno(), self.nbytes)
        self.array = np.frombuffer(self.mm, dtype=np.float32).reshape(h, w, 4)
        self.pinned = False
        if pin:
            from . import capi
            try:
                capi.host_register(self.array)
                self.pinned = True
            except Exception:  # noqa: BLE001  (pageable copies still work)
                pass
        barrier()
        if rank == 0:
            os.unlink(self.path)  # the mappings keep it alive

    def close(self):
        if self.pinned:
            from . import capi
            try:
                capi.host_unregister(self.array)
            except Exception:  # noqa: BLE001
                pass
        self.array = None
        try:
            self.mm.close()
        except BufferError:
            pass
        self.f.close()
