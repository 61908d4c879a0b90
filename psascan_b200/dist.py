"""Multi-GPU driver: one process per GPU (torch.distributed), the block pipeline of
csrc/dist.cu with its collectives served by NCCL (or gloo on CPU buffers in the host-logic tests).

Sharding: block sorts on separate GPUs, the tail walk of every block split into
`world` contiguous shares, per-GPU gap counts summed (all-reduce over NVLink), gt bits OR-ed, and the
merge sharded by output range: every rank keeps n/world entries of the merged sequence and fetches the
range of the previous level it needs with one all-to-all.
"""
import ctypes as C
import numpy as np

COLL_BCAST, COLL_ALLREDUCE_SUM_U32, COLL_ALLTOALLV = 1, 2, 3
COLLECTIVE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_int)


class AllToAllVArgs(C.Structure):
    """psc_alltoallv_args (include/psascan_cuda.h): host struct the library passes for COLL_ALLTOALLV"""
    _fields_ = [("send_d", C.c_void_p), ("recv_d", C.c_void_p),
                ("send_bytes", C.POINTER(C.c_size_t)), ("recv_bytes", C.POINTER(C.c_size_t))]


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


def tail_share(t, rank, world):
    """contiguous share [lo, hi) of a tail of t symbols, boundaries on multiples of 32"""
    per = -(-(-(-t // world)) // 32) * 32
    return min(t, per * rank), min(t, per * (rank + 1))


class _CudaView:
    def __init__(self, ptr, nbytes, typestr):
        item = int(typestr[-1])
        self.__cuda_array_interface__ = {"shape": (nbytes // item,), "typestr": typestr, "data": (ptr, False),
                                         "version": 3, "strides": None}


def tensor_from_pointer(ptr, nbytes, dtype_str, cuda):
    """zero-copy torch view of a raw buffer: dtype_str 'u1' or 'i4'"""
    import torch
    if cuda:
        return torch.as_tensor(_CudaView(ptr, nbytes, "|u1" if dtype_str == "u1" else "<i4"), device="cuda")
    ctype = C.c_uint8 if dtype_str == "u1" else C.c_int32
    item = 1 if dtype_str == "u1" else 4
    arr = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ctype)), shape=(nbytes // item,))
    return torch.from_numpy(arr)


def make_collective(cuda=True, group=None):
    """the psc_collective_fn served by torch.distributed; returns (ctypes callback, call counter)"""
    import torch
    import torch.distributed as dist
    counter = {"bcast": 0, "allreduce": 0, "alltoall": 0, "bytes": 0}
    world = dist.get_world_size(group) if dist.is_initialized() else 1

    def fn(user, op, buf, nbytes, root):
        try:
            if op == COLL_BCAST:
                t = tensor_from_pointer(buf, nbytes, "u1", cuda)
                dist.broadcast(t, src=root, group=group)
                counter["bcast"] += 1
            elif op == COLL_ALLREDUCE_SUM_U32:
                # uint32 counters summed as int32: two's-complement addition is the same
                t = tensor_from_pointer(buf, nbytes, "i4", cuda)
                dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
                counter["allreduce"] += 1
            elif op == COLL_ALLTOALLV:
                a = C.cast(buf, C.POINTER(AllToAllVArgs)).contents
                sb = [int(a.send_bytes[h]) for h in range(world)]
                rb = [int(a.recv_bytes[h]) for h in range(world)]
                empty = torch.empty(0, dtype=torch.uint8, device="cuda" if cuda else "cpu")
                src = tensor_from_pointer(a.send_d, sum(sb), "u1", cuda) if sum(sb) else empty
                dst = tensor_from_pointer(a.recv_d, sum(rb), "u1", cuda) if sum(rb) else empty
                dist.all_to_all_single(dst, src, output_split_sizes=rb, input_split_sizes=sb, group=group)
                counter["alltoall"] += 1
            else:
                return 1
            counter["bytes"] += nbytes
            if cuda:
                torch.cuda.current_stream().synchronize()
            return 0
        except Exception as e:  # never let an exception cross the C boundary
            print("psascan_b200.dist collective failed:", repr(e), flush=True)
            return 2

    return COLLECTIVE_FN(fn), counter


def _rank_world(rank, world):
    import torch.distributed as dist
    if world is None:
        world = dist.get_world_size() if dist.is_initialized() else 1
        rank = dist.get_rank() if dist.is_initialized() else 0
    return rank, world


def construct_sa_distributed(ctx, text_dev_ptr, n, out_dev_ptr, max_block, rank=None, world=None):
    """SPMD call: every rank passes the whole text (device pointer) and gets the whole SA."""
    rank, world = _rank_world(rank, world)
    cb, counter = make_collective(cuda=True)
    f = ctx.lib.psc_construct_sa_dist_d
    rc = f(ctx.h, C.c_void_p(text_dev_ptr), C.c_long(n), C.c_long(max_block), C.c_void_p(out_dev_ptr),
           C.c_int(rank), C.c_int(world), cb, None)
    ctx._chk(rc)
    return counter


def slice_capacity(ctx, n, world):
    """entries the slice buffer of construct_sa_slices must hold"""
    ctx.lib.psc_dist_slice_capacity.restype = C.c_long
    return int(ctx.lib.psc_dist_slice_capacity(C.c_long(n), C.c_int(world)))


def construct_sa_slices(ctx, text_dev_ptr, n, out_slice_dev_ptr, max_block, rank=None, world=None):
    """SPMD call: every rank passes the whole text and keeps its slice of the SA (5-byte entries
    [beg, beg + length) in out_slice_dev_ptr); returns (beg, length, collective counters)."""
    rank, world = _rank_world(rank, world)
    cb, counter = make_collective(cuda=True)
    beg, length = C.c_long(0), C.c_long(0)
    rc = ctx.lib.psc_construct_sa_dist_slice_d(ctx.h, C.c_void_p(text_dev_ptr), C.c_long(n), C.c_long(max_block),
                                               C.c_void_p(out_slice_dev_ptr), C.byref(beg), C.byref(length),
                                               C.c_int(rank), C.c_int(world), cb, None)
    ctx._chk(rc)
    return beg.value, length.value, counter
