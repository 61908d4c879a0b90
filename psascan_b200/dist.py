"""Multi-GPU driver: one process per GPU (torchrun).  The schedule and every collective live in the
library (csrc/dist.cu over NCCL, csrc/comm.cu); torch.distributed is only used to hand the NCCL id of
rank 0 to the other ranks.  The plan helpers below restate, in plain Python, how csrc/dist.cu cuts the
work (block owners, tail shares, gap slices, output shares, the split of an output boundary over the
blocks); tests/test_dist_gloo.py runs them with world_size 2 on CPU.
"""
import ctypes as C
import numpy as np

COARSE_TILE = 2048


def share_begin(g, length, world):
    """first output slot of rank g when `length` slots are shared out (csrc/dist.cu share_begin)"""
    per = (-(-length // world) + 3) // 4 * 4
    return min(length, per * g)


def block_bounds(n, max_block):
    """[(beg, end)] left to right, as partial_sufsort.hpp:564-570 lays blocks out."""
    nb = -(-n // max_block)
    return [(i * max_block, min(n, (i + 1) * max_block)) for i in range(nb)]


def block_owner(block_id, n_blocks, world):
    """rightmost block first: rank r owns blocks B-1-r, B-1-r-world, ... (csrc/dist.cu)"""
    return (n_blocks - 1 - block_id) % world


def tail_shares(end, n, world):
    """(first gt word touched, words per rank, world + 1 tail-relative boundaries): rank g walks tail
    positions [lo[g], lo[g+1]) of the block ending at `end` and writes the gt bits [end + lo[g], end +
    lo[g+1]), whole 32-bit words except at both ends (csrc/dist.cu tail_shares)"""
    wl, wh = end >> 5, (n - 1) >> 5
    per_w = -(-(wh - wl + 1) // world)
    lo = [min(n, max(end, 32 * (wl + g * per_w))) - end for g in range(world + 1)]
    return wl, per_w, lo


def gap_slice_len(m, world):
    """entries of a block's gap array (m + 1 counters) every rank keeps after the reduce-scatter"""
    return (-(-(m + 1) // world) + 3) // 4 * 4


def split_in_slice(gap_slice, u_off, e_base, last_slice, q):
    """merge_split_kernel (csrc/finalmerge.cu) for one query: the slice holds gap[u_off : u_off + cnt),
    E[u_off] = e_base.  Returns (u, E[u]) for the largest u with u + E[u] <= q when that u lies in the
    slice, else (0, 0): the answers of all slices of a block are summed."""
    cnt = len(gap_slice)
    if cnt == 0:
        return 0, 0
    total = int(np.sum(gap_slice, dtype=np.uint64))
    if not (u_off + e_base <= q):
        return 0, 0
    if not last_slice and (u_off + cnt + e_base + total <= q):
        return 0, 0
    k, e = 0, e_base
    while k < cnt - 1 and (u_off + k + 1) + e + int(gap_slice[k]) <= q:
        e += int(gap_slice[k])
        k += 1
    return u_off + k, e


def merge_boundaries(n, world, chunk):
    """(all chunk boundaries in order, index of every rank's first one): rank g's output range cut into
    chunks (csrc/dist.cu dist_merge)"""
    bd, first = [], []
    for g in range(world):
        first.append(len(bd))
        bd.extend(range(share_begin(g, n, world), share_begin(g + 1, n, world), chunk))
    first.append(len(bd))
    bd.append(n)
    return bd, first


def cascade_chunk(levels):
    """finalmerge.cu cascade_chunk in numpy: levels = [(sa_slice + beg as int64, gap_slice, skip)] left to
    right; returns the merged chunk"""
    merged = np.zeros(0, dtype=np.int64)
    for sa, gap, skip in reversed(levels):
        cb = len(sa)
        if cb == 0:
            continue
        if len(merged) == 0:
            merged = sa.copy()
            continue
        g = np.asarray(gap[:cb], dtype=np.int64).copy()
        g[0] -= skip
        slot = np.arange(cb, dtype=np.int64) + np.cumsum(g)
        out = np.zeros(cb + len(merged), dtype=np.int64)
        mask = np.zeros(len(out), dtype=bool)
        mask[slot] = True
        out[mask] = sa
        out[~mask] = merged
        merged = out
    return merged


# ---- the library over NCCL -------------------------------------------------------------------------------
def init(ctx, rank=None, world=None):
    """collective: joins `ctx` to an NCCL communicator of all torch.distributed ranks (psc_dist_init)"""
    import torch.distributed as dist
    if world is None:
        world = dist.get_world_size() if dist.is_initialized() else 1
        rank = dist.get_rank() if dist.is_initialized() else 0
    ident = (C.c_uint8 * 128)()
    if world > 1:
        box = [None]
        if rank == 0:
            ctx._chk(ctx.lib.psc_dist_unique_id(ident))
            box[0] = bytes(ident)
        dist.broadcast_object_list(box, src=0)
        ident = (C.c_uint8 * 128).from_buffer_copy(box[0])
    ctx._chk(ctx.lib.psc_dist_init(ctx.h, ident, C.c_int(rank), C.c_int(world)))
    ctx.dist_rank, ctx.dist_world = rank, world
    return rank, world


def finalize(ctx):
    ctx.lib.psc_dist_finalize(ctx.h)


def comm_stats(ctx):
    calls, nbytes = C.c_long(0), C.c_double(0)
    ctx._chk(ctx.lib.psc_dist_stats(ctx.h, C.byref(calls), C.byref(nbytes)))
    return {"collectives": calls.value, "bytes": nbytes.value}


def slice_capacity(ctx, n, world):
    """entries the slice buffer of construct_sa_slices must hold"""
    ctx.lib.psc_dist_slice_capacity.restype = C.c_long
    return int(ctx.lib.psc_dist_slice_capacity(C.c_long(n), C.c_int(world)))


def construct_sa_slices(ctx, text_dev_ptr, n, out_slice_dev_ptr, max_block):
    """SPMD call after init(): every rank passes the whole text and keeps its slice of the SA (5-byte entries
    [beg, beg + length) in out_slice_dev_ptr); returns (beg, length, collective counters)."""
    beg, length = C.c_long(0), C.c_long(0)
    rc = ctx.lib.psc_construct_sa_dist_slice_d(ctx.h, C.c_void_p(text_dev_ptr), C.c_long(n), C.c_long(max_block),
                                               C.c_void_p(out_slice_dev_ptr), C.byref(beg), C.byref(length))
    ctx._chk(rc)
    return beg.value, length.value, (comm_stats(ctx) if getattr(ctx, "dist_world", 1) > 1 else {})


def construct_sa_distributed(ctx, text_dev_ptr, n, out_dev_ptr, max_block):
    """SPMD call after init(): every rank passes the whole text (device pointer) and gets the whole SA."""
    rc = ctx.lib.psc_construct_sa_dist_d(ctx.h, C.c_void_p(text_dev_ptr), C.c_long(n), C.c_long(max_block),
                                         C.c_void_p(out_dev_ptr))
    ctx._chk(rc)
    return comm_stats(ctx) if getattr(ctx, "dist_world", 1) > 1 else {}
