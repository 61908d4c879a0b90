"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: the collective adapter that serves
the library's psc_collective_fn, and the sharding schedule of csrc/dist.cu restated with the
oracle as the compute engine (blocks sorted on their owner, tail shares walked separately, gap
counts summed by all-reduce, gt bits OR-ed) -- the result must be the suffix array."""
import os
import sys
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))


def _worker(rank, world, port, q):
    import ctypes as C
    import oracle_lib as O
    from psascan_b200 import dist as D, textgen
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # 1. the collective adapter on raw host pointers
        cb, counter = D.make_collective(cuda=False)
        buf = np.full(1000, rank + 1, dtype=np.uint32)
        assert cb(None, D.COLL_ALLREDUCE_SUM_U32, buf.ctypes.data, buf.nbytes, 0) == 0
        assert (buf == sum(range(1, world + 1))).all()
        big = np.full(4, 0xF0000000, dtype=np.uint32)   # wraps like uint32 addition
        cb(None, D.COLL_ALLREDUCE_SUM_U32, big.ctypes.data, big.nbytes, 0)
        assert (big == np.uint32((0xF0000000 * world) & 0xFFFFFFFF)).all()
        b = np.arange(77, dtype=np.uint8) if rank == 1 else np.zeros(77, dtype=np.uint8)
        assert cb(None, D.COLL_BCAST, b.ctypes.data, b.nbytes, 1) == 0
        assert (b == np.arange(77, dtype=np.uint8)).all()
        assert counter["allreduce"] == 2 and counter["bcast"] == 1
        # all-to-all with ragged shares: rank r sends (h + 1) * (r + 2) bytes of value 10 * r + h to rank h
        sb = [(h + 1) * (rank + 2) for h in range(world)]
        rb = [(rank + 1) * (h + 2) for h in range(world)]
        src = np.concatenate([np.full(sb[h], 10 * rank + h, dtype=np.uint8) for h in range(world)])
        dst = np.zeros(sum(rb), dtype=np.uint8)
        sb_c, rb_c = (C.c_size_t * world)(*sb), (C.c_size_t * world)(*rb)
        args = D.AllToAllVArgs(src.ctypes.data, dst.ctypes.data, sb_c, rb_c)
        assert cb(None, D.COLL_ALLTOALLV, C.addressof(args), dst.nbytes, 0) == 0
        assert np.array_equal(dst, np.concatenate([np.full(rb[h], 10 * h + rank, dtype=np.uint8) for h in range(world)]))
        assert counter["alltoall"] == 1

        # 2. the distributed schedule with the oracle as the engine
        for gen, kw, n, mb in [("dna", dict(seed=3), 5000, 1300), ("periodic", dict(period=37, sigma=3), 4000, 999),
                               ("random", dict(seed=5), 3000, 3000)]:
            text = textgen.GENERATORS[gen](n, **kw)
            blocks = D.block_bounds(n, mb)
            B = len(blocks)
            z = np.zeros(n + 2, dtype=np.uint8)
            # phase 1: my blocks (gt_end by plain comparison = the true order, here from the full SA)
            ref = O.sa(text)
            isa = np.empty(n + 1, dtype=np.int64); isa[ref] = np.arange(n); isa[n] = -1
            mine = {}
            for bid, (beg, end) in enumerate(blocks):
                if D.block_owner(bid, B, world) != rank:
                    continue
                gt_end = z.copy(); gt_end[:n + 1] = isa[:n + 1] > isa[end] if end < n else 0
                mine[bid] = O.sort_block(text, beg, end, gt_end)
            # phase 2
            gt_cur = z.copy()
            pieces = []
            shard = np.zeros(0, dtype=np.int64)   # my slice of the merged sequence (csrc/dist.cu, sharded merge)
            right_len = 0
            for bid in range(B - 1, -1, -1):
                beg, end = blocks[bid]
                m, t = end - beg, n - end
                own = D.block_owner(bid, B, world)
                if own == rank:
                    psa, bwt, gtb, i0 = mine[bid]
                    pack = [torch.from_numpy(psa.copy()), torch.from_numpy(bwt.copy()), torch.from_numpy(gtb.copy()),
                            torch.tensor([i0])]
                else:
                    pack = [torch.zeros(m, dtype=torch.int64), torch.zeros(m, dtype=torch.uint8),
                            torch.zeros(n + 2, dtype=torch.uint8), torch.zeros(1, dtype=torch.int64)]
                for x in pack:
                    dist.broadcast(x, src=own)
                psa, bwt, gtb, i0 = pack[0].numpy(), pack[1].numpy(), pack[2].numpy(), int(pack[3][0])
                gap = np.zeros(m + 1, dtype=np.int64)
                gt_new = z.copy()
                if t > 0:
                    lo, hi = D.tail_share(t, rank, world)
                    if hi > lo:
                        start = O.initial_rank(text, beg, end, psa, end + hi, end, gt_cur)
                        g, gout, _ = O.compute_gap(text, bwt, i0, text[end - 1], end + lo, end + hi, start, gt_cur)
                        gap += g
                        gt_new[end + lo + 1:end + hi + 1] = gout[end + lo + 1:end + hi + 1]
                    gt32 = torch.from_numpy(gap.astype(np.int32))
                    dist.all_reduce(gt32)
                    gap = gt32.numpy().astype(np.int64)
                    gw = torch.from_numpy(gt_new.astype(np.int32))
                    dist.all_reduce(gw)
                    gt_new = gw.numpy().astype(np.uint8)
                    assert gap.sum() == t
                gt_new[beg + 1:end + 1] |= gtb[beg + 1:end + 1]
                pieces.append((beg, m, psa, gap))
                gt_cur = gt_new
                # one level of the merge sharded by output range, with the exchange plan of csrc/dist.cu
                R, L = right_len, right_len + m
                a, b = D.share_begin(rank, L, world), D.share_begin(rank + 1, L, world)
                if R == 0:
                    shard = beg + psa[a:b].astype(np.int64)
                else:
                    E = np.concatenate([[0], np.cumsum(gap)])[:m + 1]          # exclusive scan
                    A = np.arange(m + 1) + E                                     # run u starts at slot A[u]
                    Q = [D.share_begin(g, L, world) for g in range(world + 1)]
                    cnt = [int(np.searchsorted(A, q, side="right")) - 1 for q in Q]
                    ra = [q - c for q, c in zip(Q, cnt)]
                    P = [D.share_begin(g, R, world) for g in range(world + 1)]
                    ov = lambda a0, a1, b0, b1: max(0, min(a1, b1) - max(a0, b0))
                    sb = [8 * ov(P[rank], P[rank + 1], ra[h], ra[h + 1]) for h in range(world)]
                    rb = [8 * ov(P[h], P[h + 1], ra[rank], ra[rank + 1]) for h in range(world)]
                    assert ra[0] == 0 and ra[world] == R and sum(sb) == 8 * len(shard)
                    recv = np.zeros(sum(rb) // 8, dtype=np.int64)
                    src = np.ascontiguousarray(shard)
                    args = D.AllToAllVArgs(src.ctypes.data, recv.ctypes.data, (C.c_size_t * world)(*sb),
                                           (C.c_size_t * world)(*rb))
                    assert cb(None, D.COLL_ALLTOALLV, C.addressof(args), recv.nbytes, 0) == 0
                    out = np.empty(b - a, dtype=np.int64)
                    is_blk = np.zeros(b - a, dtype=bool)
                    slot = np.arange(m) + E[1:m + 1]                             # block suffix u sits at u + E[u+1]
                    sel = (slot >= a) & (slot < b)
                    is_blk[slot[sel] - a] = True
                    out[is_blk] = beg + psa[sel].astype(np.int64)
                    assert (~is_blk).sum() == len(recv)
                    out[~is_blk] = recv
                    shard = out
                right_len = L
            pieces.sort()
            sa = O.merge([p[0] for p in pieces], [p[1] for p in pieces], [p[2] for p in pieces], [p[3] for p in pieces])
            assert np.array_equal(sa, ref), (gen, rank)
            parts = [None] * world
            dist.all_gather_object(parts, shard)
            assert np.array_equal(np.concatenate(parts), ref), (gen, rank, "sharded merge")
        q.put((rank, "ok"))
    except Exception as e:
        import traceback
        q.put((rank, "FAIL " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_schedule_and_collectives(world):
    """world 2, and 3 for shares that do not divide evenly (ragged all-to-all splits, empty overlaps)"""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + world
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(r[1] == "ok" for r in res), res


def test_plan_helpers():
    from psascan_b200 import dist as D
    assert D.block_bounds(10, 4) == [(0, 4), (4, 8), (8, 10)]
    assert [D.block_owner(b, 5, 2) for b in range(5)] == [0, 1, 0, 1, 0]
    for t in (1, 31, 32, 33, 1000, 4097):
        for w in (1, 2, 3, 8):
            shares = [D.tail_share(t, r, w) for r in range(w)]
            assert shares[0][0] == 0 and shares[-1][1] == t
            for a, b in zip(shares, shares[1:]):
                assert a[1] == b[0] and a[1] % 32 == 0 or a[1] == t


@pytest.mark.parametrize("world", [1, 2, 3, 5, 8])
def test_sharded_merge_plan_simulated(world):
    """The exchange plan of the sharded merge (csrc/dist.cu: share boundaries, block suffixes before each
    boundary, who sends which part of its share to whom) simulated for all ranks in one process, on random
    gap arrays including empty shares (fewer slots than ranks) and lopsided interleavings."""
    from psascan_b200 import dist as D
    rng = np.random.default_rng(world)
    ov = lambda a0, a1, b0, b1: max(0, min(a1, b1) - max(a0, b0))
    for trial in range(30):
        nblocks = int(rng.integers(2, 6))
        sizes = [int(rng.integers(1, 40)) for _ in range(nblocks)]
        # a random "suffix array": global order of all positions; block b owns positions [beg_b, beg_b + m_b)
        n = sum(sizes)
        order = rng.permutation(n) if trial % 3 else np.arange(n)          # also the all-left-first extreme
        begs = np.concatenate([[0], np.cumsum(sizes)])[:-1]
        rank_of = np.empty(n, dtype=np.int64); rank_of[order] = np.arange(n)
        shards = [np.zeros(0, dtype=np.int64) for _ in range(world)]
        R = 0
        for b in range(nblocks - 1, -1, -1):
            m, beg = sizes[b], int(begs[b])
            mine = np.arange(beg, beg + m)
            psa = mine[np.argsort(rank_of[mine])] - beg                    # block-relative SA
            right = np.arange(beg + m, n)
            # gap[u] = number of right-side suffixes between block suffixes u-1 and u
            rr = np.sort(rank_of[right]); br = rank_of[psa + beg]
            gap = np.diff(np.concatenate([[0], np.searchsorted(rr, br), [len(rr)]]))
            L = R + m
            Q = [D.share_begin(g, L, world) for g in range(world + 1)]
            if R == 0:
                shards = [beg + psa[Q[g]:Q[g + 1]] for g in range(world)]
            else:
                E = np.concatenate([[0], np.cumsum(gap)])[:m + 1]
                A = np.arange(m + 1) + E
                cnt = [int(np.searchsorted(A, q, side="right")) - 1 for q in Q]
                ra = [q - c for q, c in zip(Q, cnt)]
                P = [D.share_begin(g, R, world) for g in range(world + 1)]
                assert ra[0] == 0 and ra[world] == R and all(x <= y for x, y in zip(ra, ra[1:]))
                new = []
                slot = np.arange(m) + E[1:m + 1]
                for g in range(world):
                    recv = np.concatenate([shards[h][max(P[h], ra[g]) - P[h]:max(P[h], ra[g]) - P[h] + ov(P[h], P[h + 1], ra[g], ra[g + 1])]
                                           for h in range(world)]) if world else np.zeros(0, dtype=np.int64)
                    a, bnd = Q[g], Q[g + 1]
                    out = np.empty(bnd - a, dtype=np.int64)
                    is_blk = np.zeros(bnd - a, dtype=bool)
                    sel = (slot >= a) & (slot < bnd)
                    is_blk[slot[sel] - a] = True
                    out[is_blk] = beg + psa[sel]
                    assert (~is_blk).sum() == len(recv) == ra[g + 1] - ra[g]
                    out[~is_blk] = recv
                    new.append(out)
                shards = new
            R = L
            assert [len(x) for x in shards] == [Q[g + 1] - Q[g] for g in range(world)]
        merged = np.concatenate(shards)
        assert np.array_equal(merged, np.argsort(rank_of)), (world, trial)
