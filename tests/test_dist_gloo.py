"""CPU tests of the multi-GPU host logic (csrc/dist.cu restated in psascan_b200/dist.py):
  * world_size 2 and 3 over gloo, with the oracle as the compute engine: blocks sorted on their owner, the
    owner hands out start ranks + BWT (never the block SA), tail shares walked separately, gap counts
    reduce-scattered into slices, gt bits gathered by word range, boundaries split by a distributed search
    over the gap slices, ONE exchange of SA/gap slices, merge by output range -- the result must be the
    suffix array;
  * the split + exchange + cascade plan simulated for all ranks in one process on random gap arrays."""
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


def split_all_levels(D, gaps, sizes, world, boundaries):
    """the distributed split of csrc/dist.cu dist_merge for all ranks at once: gaps[k] = gap array of block k
    (m_k + 1 counters; None for the rightmost block).  Returns Q, U, E as [level][boundary] lists."""
    B, nb = len(sizes), len(boundaries)
    Q, U, E = [list(boundaries)], [], []
    for k in range(B):
        if k == B - 1:
            U.append(list(Q[k])); E.append([0] * nb)
            break
        m, per = sizes[k], D.gap_slice_len(sizes[k], world)
        u_sum, e_sum = [0] * nb, [0] * nb
        e_base = 0
        for h in range(world):
            sl = gaps[k][per * h:min(m + 1, per * (h + 1))]
            last = len(sl) > 0 and per * h + len(sl) == m + 1
            for i, q in enumerate(Q[k]):
                u, e = D.split_in_slice(sl, per * h, e_base, last, q)
                u_sum[i] += u; e_sum[i] += e
            e_base += int(np.sum(sl, dtype=np.int64))
        U.append(u_sum); E.append(e_sum)
        Q.append([q - u for q, u in zip(Q[k], u_sum)])
    return Q, U, E


@pytest.mark.parametrize("world", [1, 2, 3, 5, 8])
def test_split_and_cascade_simulated(world):
    """random block sizes and interleavings (also the all-left-first extreme and shares with no slots):
    every output chunk merged from the slices the plan names must reproduce the global order"""
    from psascan_b200 import dist as D
    rng = np.random.default_rng(world)
    for trial in range(30):
        nblocks = int(rng.integers(2, 6))
        sizes = [int(rng.integers(1, 40)) for _ in range(nblocks)]
        n = sum(sizes)
        order = rng.permutation(n) if trial % 3 else np.arange(n)
        begs = [int(x) for x in np.concatenate([[0], np.cumsum(sizes)])[:-1]]
        rank_of = np.empty(n, dtype=np.int64); rank_of[order] = np.arange(n)
        psas, gaps = [], []
        for b in range(nblocks):
            m, beg = sizes[b], begs[b]
            mine = np.arange(beg, beg + m)
            psa = mine[np.argsort(rank_of[mine])] - beg
            right = np.arange(beg + m, n)
            rr = np.sort(rank_of[right]); br = rank_of[psa + beg]
            gap = np.diff(np.concatenate([[0], np.searchsorted(rr, br), [len(rr)]]))
            psas.append(psa); gaps.append(gap if b < nblocks - 1 else None)
        chunk = int(rng.choice([4, 8, 20, 1000]))
        bd, first = D.merge_boundaries(n, world, chunk)
        assert bd[0] == 0 and bd[-1] == n and all(x < y for x, y in zip(bd, bd[1:])) or n == 0
        Q, U, E = split_all_levels(D, gaps, sizes, world, bd)
        for k in range(nblocks):   # boundaries split monotonically, all of every block is handed out
            assert U[k][0] == 0 and U[k][-1] == sizes[k] and all(x <= y for x, y in zip(U[k], U[k][1:]))
        out = []
        for i in range(len(bd) - 1):
            levels = []
            for k in range(nblocks):
                ua, ub = U[k][i], U[k][i + 1]
                skip = 0 if k == nblocks - 1 else Q[k][i] - (ua + E[k][i])
                g = gaps[k][ua:ub + 1] if k < nblocks - 1 else None
                levels.append((begs[k] + psas[k][ua:ub].astype(np.int64), g, skip))
            got = D.cascade_chunk(levels)
            assert len(got) == bd[i + 1] - bd[i]
            out.append(got)
        assert np.array_equal(np.concatenate(out), np.argsort(rank_of)), (world, trial)


def test_plan_helpers():
    from psascan_b200 import dist as D
    assert D.block_bounds(10, 4) == [(0, 4), (4, 8), (8, 10)]
    assert [D.block_owner(b, 5, 2) for b in range(5)] == [0, 1, 0, 1, 0]
    for n, end in ((100, 1), (100, 31), (100, 32), (1000, 33), (4097, 1000), (5000, 4999), (64, 32)):
        for w in (1, 2, 3, 8):
            wl, per_w, lo = D.tail_shares(end, n, w)
            assert lo[0] == 0 and lo[-1] == n - end and all(a <= b for a, b in zip(lo, lo[1:]))
            for g in range(1, w):   # interior boundaries sit on word boundaries of the gt bit array
                assert (end + lo[g]) % 32 == 0 or lo[g] == n - end
            for g in range(w):      # rank g's bits lie inside its word range
                if lo[g + 1] > lo[g]:
                    assert (end + lo[g]) >> 5 >= wl + g * per_w and (end + lo[g + 1] - 1) >> 5 < wl + (g + 1) * per_w
    for m in (1, 3, 4, 1000):
        for w in (1, 2, 3, 8):
            per = D.gap_slice_len(m, w)
            assert per % 4 == 0 and per * w >= m + 1


def _worker(rank, world, port, q):
    import oracle_lib as O
    from psascan_b200 import dist as D, textgen
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        for gen, kw, n, mb in [("dna", dict(seed=3), 5000, 1300), ("periodic", dict(period=37, sigma=3), 4000, 999),
                               ("random", dict(seed=5), 3000, 3000), ("markov", dict(seed=9), 2100, 300)]:
            text = textgen.GENERATORS[gen](n, **kw)
            blocks = D.block_bounds(n, mb)
            B = len(blocks)
            sizes = [e - b for b, e in blocks]
            z = np.zeros(n + 2, dtype=np.uint8)
            ref = O.sa(text)
            isa = np.empty(n + 1, dtype=np.int64); isa[ref] = np.arange(n); isa[n] = -1
            # phase 1: my blocks (gt_end by plain comparison = the true order, here read off the full SA)
            mine = {}
            for bid, (beg, end) in enumerate(blocks):
                if D.block_owner(bid, B, world) != rank:
                    continue
                gt_end = z.copy(); gt_end[:n + 1] = isa[:n + 1] > isa[end] if end < n else 0
                mine[bid] = O.sort_block(text, beg, end, gt_end)
            # phase 2: right to left
            gt_cur = z.copy()
            my_sa, my_gap = {}, {}
            for bid in range(B - 1, -1, -1):
                beg, end = blocks[bid]
                m, t = end - beg, n - end
                own = D.block_owner(bid, B, world)
                wl, per_w, lo = D.tail_shares(end, n, world) if t > 0 else (0, 0, [0] * (world + 1))
                if own == rank:
                    psa, bwt, gtb, i0 = mine[bid]
                    my_sa[bid] = psa
                    # the owner computes the start rank of every rank's walker (K5 needs the block SA)
                    starts = [O.initial_rank(text, beg, end, psa, end + lo[g + 1], end, gt_cur) if t > 0 and lo[g + 1] > lo[g] else 0
                              for g in range(world)]
                    pack = [torch.from_numpy(bwt.copy()), torch.from_numpy(gtb.copy()), torch.tensor([i0] + starts)]
                else:
                    pack = [torch.zeros(m, dtype=torch.uint8), torch.zeros(n + 2, dtype=torch.uint8),
                            torch.zeros(1 + world, dtype=torch.int64)]
                for x in pack:   # BWT, gt_begin bits, i0 and start ranks -- NOT the block SA
                    dist.broadcast(x, src=own)
                bwt, gtb, i0, starts = pack[0].numpy(), pack[1].numpy(), int(pack[2][0]), [int(v) for v in pack[2][1:]]
                gt_new = z.copy()
                if t > 0:
                    per = D.gap_slice_len(m, world)
                    gap_full = np.zeros(per * world, dtype=np.int64)
                    mybits = np.zeros(per_w * 32, dtype=np.uint8)     # my word range of the new gt bits
                    if lo[rank + 1] > lo[rank]:
                        g, gout, _ = O.compute_gap(text, bwt, i0, text[end - 1], end + lo[rank], end + lo[rank + 1],
                                                   starts[rank], gt_cur)
                        gap_full[:m + 1] += g
                        for j in range(end + lo[rank] + 1, end + lo[rank + 1] + 1):   # bit index = position - 1
                            mybits[(j - 1) - 32 * (wl + rank * per_w)] = gout[j]
                    # reduce-scatter (gloo has none: all-reduce, keep my slice)
                    gsum = torch.from_numpy(gap_full.copy())
                    dist.all_reduce(gsum)
                    assert int(gsum.sum()) == t
                    my_gap[bid] = gsum.numpy()[per * rank:min(m + 1, per * (rank + 1))].copy()
                    # all-gather of disjoint word ranges
                    allbits = [torch.zeros(per_w * 32, dtype=torch.uint8) for _ in range(world)]
                    dist.all_gather(allbits, torch.from_numpy(mybits))
                    for g in range(world):
                        base = 32 * (wl + g * per_w)
                        for r, v in enumerate(allbits[g].numpy()):
                            if v and base + r + 1 <= n:
                                gt_new[base + r + 1] = 1
                gt_new[beg + 1:end + 1] |= gtb[beg + 1:end + 1]
                gt_cur = gt_new
            # phase 3: distributed split (one all-reduce per level), one exchange, cascade per chunk
            bd, first = D.merge_boundaries(n, world, 100)
            nb = len(bd)
            Q, U, E = [list(bd)], [], []
            for k in range(B):
                if k == B - 1:
                    U.append(list(Q[k])); E.append([0] * nb)
                    break
                m, per = sizes[k], D.gap_slice_len(sizes[k], world)
                tots = [None] * world
                dist.all_gather_object(tots, int(my_gap[k].sum()))
                e_base = sum(tots[:rank])
                last = len(my_gap[k]) > 0 and per * rank + len(my_gap[k]) == m + 1
                ue = torch.tensor([D.split_in_slice(my_gap[k], per * rank, e_base, last, qq) for qq in Q[k]], dtype=torch.int64)
                dist.all_reduce(ue)
                U.append([int(v) for v in ue[:, 0]]); E.append([int(v) for v in ue[:, 1]])
                Q.append([a - b for a, b in zip(Q[k], U[k])])
            need = lambda k, g: (U[k][first[g]], U[k][first[g + 1]])
            msgs = []   # (block, kind, to, first entry, data): what the NCCL send/recv group carries
            for k in range(B):
                for g in range(world):
                    lo_, hi_ = need(k, g)
                    if D.block_owner(k, B, world) == rank and hi_ > lo_:
                        msgs.append((k, "sa", g, lo_, my_sa[k][lo_:hi_]))
                    if k < B - 1:
                        per = D.gap_slice_len(sizes[k], world)
                        s0, s1 = per * rank, per * rank + len(my_gap[k])
                        a, b = max(s0, lo_), min(s1, hi_)
                        if b > a:
                            msgs.append((k, "gap", g, a, my_gap[k][a - s0:b - s0]))
            allm = [None] * world
            dist.all_gather_object(allm, msgs)
            rsa = {k: np.zeros(need(k, rank)[1] - need(k, rank)[0], dtype=np.int64) for k in range(B)}
            rgap = {k: np.zeros(need(k, rank)[1] - need(k, rank)[0] + 1, dtype=np.int64) for k in range(B)}
            filled = {k: 0 for k in range(B)}
            for ms in allm:
                for k, kind, to, at, data in ms:
                    if to != rank:
                        continue
                    d = at - need(k, rank)[0]
                    if kind == "sa":
                        rsa[k][d:d + len(data)] = data
                    else:
                        rgap[k][d:d + len(data)] = data
                        filled[k] += len(data)
            for k in range(B - 1):   # the gap slices of all holders tile my range exactly
                assert filled[k] == need(k, rank)[1] - need(k, rank)[0]
            out = []
            for i in range(first[rank], first[rank + 1]):
                levels = []
                for k in range(B):
                    ua, ub = U[k][i], U[k][i + 1]
                    d = ua - need(k, rank)[0]
                    skip = 0 if k == B - 1 else Q[k][i] - (ua + E[k][i])
                    levels.append((blocks[k][0] + rsa[k][d:d + ub - ua], rgap[k][d:d + ub - ua + 1], skip))
                out.append(D.cascade_chunk(levels))
            shard = np.concatenate(out) if out else np.zeros(0, dtype=np.int64)
            parts = [None] * world
            dist.all_gather_object(parts, shard)
            assert np.array_equal(np.concatenate(parts), ref), (gen, rank, "merge by output range")
        q.put((rank, "ok"))
    except Exception:
        import traceback
        q.put((rank, "FAIL " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_schedule_over_gloo(world):
    """world 2, and 3 for shares and slices that do not divide evenly"""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + world
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(r[1] == "ok" for r in res), res
