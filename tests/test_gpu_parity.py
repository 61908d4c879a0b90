"""GPU parity tests (run on the B200 box): every stage of the CUDA path, called through the C ABI
(psascan_b200.lib -> libpsascan_cuda.so), against the CPU oracle on the same seeded inputs, and the
whole path against the reference's golden .sa5 files.  Integer work: bit-exact."""
import numpy as np
import pytest
import oracle_lib as O
from golden_cases import golden_cases, golden_text
from psascan_b200 import textgen
from psascan_b200.lib import decode_sa5

pytestmark = pytest.mark.gpu
CASES = golden_cases()


def rev_bits(flags_by_pos, n, lo, hi):
    return O.pack_gt_rev(flags_by_pos, n, lo, hi)


def texts_small():
    rng = np.random.default_rng(123)
    out = [("one", np.array([7], dtype=np.uint8)), ("two_eq", np.array([3, 3], dtype=np.uint8)),
           ("all_a", np.full(1000, 97, dtype=np.uint8)), ("zeros", np.zeros(777, dtype=np.uint8)),
           ("ab_period", np.tile(np.array([0, 1], dtype=np.uint8), 2500)),
           ("dna", textgen.dna(30000, seed=5)), ("rand255", textgen.random_bytes(50000, seed=6)),
           ("markov", textgen.markov(40000, seed=8)),
           ("rep", textgen.repetitive(60000, seed=4, seed_len=3000, rate=1e-3)),
           ("periodic0", textgen.periodic(20000, period=97, sigma=3)),
           ("tiny_rand", rng.integers(0, 3, size=257, dtype=np.uint8))]
    return out


@pytest.mark.parametrize("name,text", texts_small(), ids=[t[0] for t in texts_small()])
def test_construct_sa_single_block(ctx, name, text):
    got = decode_sa5(ctx.construct_sa(text))
    assert np.array_equal(got, O.sa(text))


@pytest.mark.parametrize("name,text", texts_small(), ids=[t[0] for t in texts_small()])
@pytest.mark.parametrize("nblk", [2, 3, 7])
def test_construct_sa_multi_block(ctx, name, text, nblk):
    n = len(text)
    mb = max(1, -(-n // nblk))
    got = decode_sa5(ctx.construct_sa(text, max_block=mb))
    assert np.array_equal(got, O.sa(text)), (name, mb)


@pytest.mark.parametrize("name,gen,kw,n,sa5", CASES, ids=[c[0] for c in CASES])
def test_golden_sa5_bit_exact(ctx, name, gen, kw, n, sa5):
    text = golden_text(gen, kw, n)
    assert np.array_equal(ctx.construct_sa(text), sa5)
    assert np.array_equal(ctx.construct_sa(text, max_block=n // 3 + 5), sa5)
    assert np.array_equal(ctx.construct_sa(text, max_block=1024), sa5)


@pytest.mark.parametrize("env", [dict(PSC_PACKED="0"), dict(PSC_PACKED="1"), dict(PSC_PACKED="1", PSC_PACKED_MAXBITS="13"),
                                 dict(PSC_PACKED="1", PSC_PACKED_MAXBITS="9"), dict(PSC_PACKED="1", PSC_PACKED_MAXBITS="21")],
                         ids=["pairs_keygen", "packed", "packed13", "packed9", "packed21"])
@pytest.mark.parametrize("sigma", [255, 200, 129])
def test_sort_modes_8bit_codes(ctx, monkeypatch, env, sigma):
    """Round 0 for 8-bit renamed codes: keys built from the text inside the first digit pass, as
    (key, position) pairs or packed into one word, also with a partly kept last symbol (few key bits:
    many ties go on to the later rounds).  Single block, and blocks with a tail (renaming by gt_end)."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    rng = np.random.default_rng(sigma)
    text = rng.integers(0, sigma, size=70001, dtype=np.uint8)
    text[5000:5600] = text[100:700]          # some long repeats: ties past round 0
    text[-300:] = text[40000:40300]
    want = O.sa(text)
    assert np.array_equal(decode_sa5(ctx.construct_sa(text)), want)
    assert np.array_equal(decode_sa5(ctx.construct_sa(text, max_block=30000)), want)


@pytest.mark.parametrize("rounds", ["0", "1", "6"])
def test_text_round_policy(ctx, monkeypatch, rounds):
    """After round 0 the sorter extends tied suffixes by symbols read from the text, again and again while
    few are tied, before it pays for the ISA and doubles: no text round, one, or up to six must all give the SA."""
    monkeypatch.setenv("PSC_TEXT_ROUNDS", rounds)
    for text in (textgen.markov(60000, seed=11), textgen.repetitive(50000, seed=4, seed_len=2000, rate=2e-3),
                 textgen.dna(40000, seed=12)):
        want = O.sa(text)
        assert np.array_equal(decode_sa5(ctx.construct_sa(text)), want)
        assert np.array_equal(decode_sa5(ctx.construct_sa(text, max_block=17000)), want)


@pytest.mark.parametrize("sigma", [2, 4, 16, 64])
def test_dense_round0_codes(ctx, sigma):
    """Alphabets of exactly 2^b symbols (one block, no renaming split): round 0 packs code - 1 at b bits, so the
    padding of the last few suffixes equals the smallest symbol; texts that END in runs of the smallest symbol
    (and contain longer runs of it elsewhere) are where that matters."""
    rng = np.random.default_rng(100 + sigma)
    for n, tail in ((50001, 0), (40000, 1), (40000, 7), (40000, 70), (3000, 200)):
        text = rng.integers(0, sigma, size=n, dtype=np.uint8)
        text[1000:1300] = 0
        text[20000 % n:20000 % n + 90] = 0
        if tail:
            text[-tail:] = 0
        assert len(np.unique(text)) == sigma
        assert np.array_equal(decode_sa5(ctx.construct_sa(text)), O.sa(text)), (sigma, n, tail)
        text2 = (sigma - 1 - text).astype(np.uint8)          # ends in the LARGEST symbol instead
        assert np.array_equal(decode_sa5(ctx.construct_sa(text2)), O.sa(text2)), (sigma, n, tail, "mirrored")


def test_high_byte_plane_forced(ctx, monkeypatch):
    """Texts above 4 GiB keep a ninth..fortieth bit plane through the merge and the 40-bit writer; PSC_FORCE_HI
    switches that path on for small texts (all high bytes are 0, the code path is the one of big texts)."""
    import torch
    monkeypatch.setenv("PSC_FORCE_HI", "1")
    for name, gen, kw, n, sa5 in CASES[:4]:
        text = golden_text(gen, kw, n)
        assert np.array_equal(ctx.construct_sa(text, max_block=n // 3 + 5), sa5), name
        d_text = torch.from_numpy(text).cuda()
        d_out = torch.zeros(5 * n, dtype=torch.uint8, device="cuda")
        ctx.construct_sa_dist_ptr(d_text.data_ptr(), n, d_out.data_ptr(), 1000)
        torch.cuda.synchronize()
        assert np.array_equal(d_out.cpu().numpy(), sa5), (name, "dist schedule")


def test_sentinel_255_rejected(ctx):
    import psascan_b200 as P
    text = textgen.dna(5000).copy()
    text[1234] = 255
    with pytest.raises(P.PscError) as e:
        ctx.construct_sa(text)
    assert e.value.code == -3


def test_empty_text(ctx):
    assert len(ctx.construct_sa(np.zeros(0, dtype=np.uint8))) == 0


@pytest.mark.parametrize("gen,kw", [("dna", dict(seed=21)), ("periodic", dict(period=53, sigma=3)),
                                    ("random", dict(seed=22)), ("repetitive", dict(seed=4, seed_len=500, rate=3e-3))])
def test_stage1_sort_block(ctx, gen, kw):
    """psc_sort_block vs orc_sort_block: SA, BWT (0 at i0), gt_begin (reversed layout), i0 --
    for the last block and for a block with a tail (renaming + gt_end)."""
    n = 9000
    text = textgen.GENERATORS[gen](n, **kw)
    z = np.zeros(n + 2, dtype=np.uint8)
    beg2, end2 = 6000, n
    psa2, bwt2, gtb2, i02 = O.sort_block(text, beg2, end2, z)
    sa, bwt, gt, i0 = ctx.sort_block(text[beg2:end2])
    assert i0 == i02 and np.array_equal(sa, psa2) and np.array_equal(bwt, bwt2)
    assert np.array_equal(gt, rev_bits(gtb2, n, beg2, end2))
    beg1, end1 = 3000, 6000
    psa1, bwt1, gtb1, i01 = O.sort_block(text, beg1, end1, gtb2)
    tl = min(end1 - beg1, n - end1)
    sa, bwt, gt, i0 = ctx.sort_block(text[beg1:end1], tail_prefix=text[end1:end1 + tl],
                                     tail_gt_rev=rev_bits(gtb2, n, end1, end1 + tl))
    assert i0 == i01 and np.array_equal(sa, psa1) and np.array_equal(bwt, bwt1)
    assert np.array_equal(gt, rev_bits(gtb1, n, beg1, end1))


@pytest.mark.parametrize("sigma", [1, 2, 4, 30, 255])
def test_stage3a_rank(ctx, sigma):
    rng = np.random.default_rng(sigma)
    m = 50000 + sigma
    bwt = rng.integers(0, sigma, size=m, dtype=np.uint8)
    i0 = 777
    bwt[i0] = 0
    r = ctx.rank_build(bwt, i0)
    cnt = np.bincount(np.delete(bwt, i0), minlength=256)
    assert np.array_equal(r.counts(), cnt)
    qi = np.concatenate([rng.integers(0, m + 1, size=4000), [0, 1, 223, 224, 225, i0, i0 + 1, m - 1, m, m + 5, -3]])
    qc = rng.integers(0, max(sigma, 2), size=len(qi)).astype(np.uint8)
    got = r.query(qi, qc)
    masked = bwt.astype(np.int64).copy()
    masked[i0] = -1
    for k in range(len(qi)):
        i = min(max(int(qi[k]), 0), m)
        assert got[k] == int((masked[:i] == qc[k]).sum()), (k, qi[k], qc[k])
    r.free()


def test_stage2_and_3b_gap(ctx):
    """psc_initial_ranks + psc_compute_gap vs orc_initial_rank / orc_compute_gap."""
    for gen, kw in [("dna", dict(seed=31)), ("periodic", dict(period=41, sigma=3)), ("markov", dict(seed=32))]:
        n, beg, end = 12000, 2000, 5000
        text = textgen.GENERATORS[gen](n, **kw)
        ref = O.sa(text)
        isa = np.empty(n + 1, dtype=np.int64); isa[ref] = np.arange(n); isa[n] = -1
        gt_end = np.zeros(n + 2, dtype=np.uint8)
        gt_end[:n + 1] = isa[:n + 1] > isa[end]
        psa, bwt, gtb, i0 = O.sort_block(text, beg, end, gt_end)
        last = text[end - 1]
        gap_o, gt_out_o, _ = O.compute_gap(text, bwt, i0, last, end, n, 0, gt_end)
        t = n - end
        # stage 2
        starts = np.array([0, 32, 64, 1000, 4992, t - 1, t], dtype=np.int64)
        after_gt = rev_bits(gt_end, n, end, n)
        got = ctx.initial_ranks(text[beg:end], psa.astype(np.uint32), text[end:], after_gt, starts)
        want = [O.initial_rank(text, beg, end, psa, end + int(s), end, gt_end) for s in starts]
        assert list(got) == want
        # stage 3: several walker layouts give the same gap / gt_out
        r = ctx.rank_build(bwt, i0)
        counts = np.bincount(text[beg:end], minlength=256)
        for slice_len in (t, 4096, 256, 64):
            wb = list(range(0, t, slice_len)) + [t]
            sr = ctx.initial_ranks(text[beg:end], psa.astype(np.uint32), text[end:], after_gt, np.array(wb[1:]))
            gap, gt_out = ctx.compute_gap(r, i0, last, counts, text[end:], wb, sr, after_gt)
            assert np.array_equal(gap.astype(np.int64), gap_o), (gen, slice_len)
            assert np.array_equal(gt_out, rev_bits(gt_out_o, n, end, n)), (gen, slice_len)
        r.free()


def test_stage4_merge_level(ctx):
    text = textgen.markov(20000, seed=41)
    n = len(text)
    ref = O.sa(text)
    cuts = [0, 6000, 13000, n]
    isa = np.empty(n, dtype=np.int64); isa[ref] = np.arange(n)
    right = np.zeros(0, dtype=np.uint8)
    for k in (2, 1, 0):
        beg, end = cuts[k], cuts[k + 1]
        in_blk = (ref >= beg) & (ref < end)
        psa = (ref[in_blk] - beg).astype(np.uint32)
        # gap[u] = tail suffixes between the block's (u-1)-th and u-th suffix
        tail_mask = ref >= end
        csum = np.concatenate([[0], np.cumsum(tail_mask)])
        pos_blk = np.nonzero(in_blk)[0]
        bounds = np.concatenate([[0], csum[pos_blk], [csum[-1]]])
        gap = np.diff(bounds).astype(np.uint32)
        right = ctx.merge_level(psa, gap, beg, right)
        want = ref[ref >= beg]
        assert np.array_equal(decode_sa5(right), want), k


def test_glue_two_piece_block(ctx):
    """gap -> bitvector, BWT merge and gap split vs the oracle (bwt_merge.hpp, compute_*_gap.hpp)."""
    rng = np.random.default_rng(5)
    len_l, len_r = 3000, 4100
    gap_l = np.zeros(len_l + 1, dtype=np.int64)
    np.add.at(gap_l, rng.integers(0, len_l + 1, size=len_r), 1)
    bv_o = O.gap_to_bv(gap_l, len_l, len_r)
    bv = ctx.gap_to_bitvector(gap_l, len_l, len_r)
    assert np.array_equal(np.unpackbits(bv, bitorder="little")[:len_l + len_r], bv_o)
    bl = rng.integers(0, 5, size=len_l, dtype=np.uint8)
    br = rng.integers(0, 5, size=len_r, dtype=np.uint8)
    out_o, i0_o = O.merge_bwt(bl, br, 17, 23, 4, bv_o)
    out, i0 = ctx.merge_bwt(bl, br, 17, 23, 4, bv)
    assert i0 == i0_o and np.array_equal(out, out_o)
    bg = rng.integers(0, 4, size=len_l + len_r + 1).astype(np.int64)
    lg_o, rg_o = O.split_gap(bv_o, len_l, len_r, bg)
    lg, rg = ctx.split_gap(bv, len_l, len_r, bg)
    assert np.array_equal(lg.astype(np.int64), lg_o) and np.array_equal(rg.astype(np.int64), rg_o)


def sa_is_valid(text, sa):
    """O(n) suffix-array check (the idea of extlib/libdivsufsort/lib/utils.c:161-241 sufcheck):
    permutation; first symbols non-decreasing; ties ordered by the rank of the next suffix."""
    n = len(text)
    if len(sa) != n or sa.min() < 0 or sa.max() >= n:
        return False
    isa = np.full(n + 1, -1, dtype=np.int64)
    isa[sa] = np.arange(n)
    if (isa[:n] < 0).any():
        return False
    c = text[sa].astype(np.int64)
    if (np.diff(c) < 0).any():
        return False
    nxt = isa[sa + 1]
    tie = np.diff(c) == 0
    return bool((nxt[1:][tie] > nxt[:-1][tie]).all())


@pytest.mark.parametrize("gen,n,mb", [("dna", 3_000_000, 0), ("random", 2_000_000, 0),
                                      ("markov", 1_500_000, 400_000), ("repetitive", 2_000_000, 700_001),
                                      # blocks above 2^24: range-partitioned ISA scatter and rank stream
                                      ("markov", 40_000_000, 18_000_000), ("repetitive", 20_000_000, 0)])
def test_large_property_check(ctx, gen, n, mb):
    text = textgen.GENERATORS[gen](n)
    sa = decode_sa5(ctx.construct_sa(text, max_block=mb))
    assert sa_is_valid(text, sa)
    if n <= 3_000_000:
        assert np.array_equal(sa, O.sa(text)) if n <= 2_000_000 else True


def test_dist_schedule_single_rank(ctx):
    """csrc/dist.cu with world = 1: sorts first (gt_end by plain comparison past the block end),
    then the right-to-left walk/merge phase -- must equal the streaming single-GPU pipeline."""
    import torch
    for name, gen, kw, n, sa5 in CASES:
        text = golden_text(gen, kw, n)
        d_text = torch.from_numpy(text).cuda()
        d_out = torch.zeros(5 * n, dtype=torch.uint8, device="cuda")
        for mb in (n, n // 2 + 3, 1000):
            d_out.zero_()
            ctx.construct_sa_dist_ptr(d_text.data_ptr(), n, d_out.data_ptr(), mb)
            torch.cuda.synchronize()
            assert np.array_equal(d_out.cpu().numpy(), sa5), (name, mb)


@pytest.mark.parametrize("env", [dict(PSC_WALK_MODE="1"), dict(PSC_WALK_MODE="2"), dict(PSC_WALK_MODE="3"),
                                 dict(PSC_WALK_MODE="3", PSC_WALK_CHUNK="2048", PSC_WALKERS="40"), dict(PSC_WALK_CHUNK="4096"),
                                 dict(PSC_WALKERS="7"), dict(PSC_MERGE_CHUNK="1000"), dict(PSC_BWT_SHIFT="8"),
                                 dict(PSC_MERGE_CHUNK="64", PSC_WALK_CHUNK="2048", PSC_WALK_MODE="2")],
                         ids=["walk_warp_aggregated", "walk_emit_partition", "walk_paged", "walk_paged_chunks",
                              "walk_multi_chunk", "seven_walkers",
                              "merge_small_chunks", "bwt_windowed", "all_small"])
def test_env_selected_paths_subprocess(env):
    """Paths chosen by environment variables read once per process (the warp-aggregated and the
    emit + partition walk, tails cut into several walk chunks, a handful of walkers, the final merge in
    many small output chunks) run in a child: every multi-block golden case must come out bit-exact,
    through the host-buffer entry point (overlapped copies) and the device-buffer one."""
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys, numpy as np, torch; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "import psascan_b200 as P\n"
        "from golden_cases import golden_cases, golden_text\n"
        "c = P.Context(0)\n"
        "for name, gen, kw, n, sa5 in golden_cases():\n"
        "    t = golden_text(gen, kw, n)\n"
        "    dt = torch.from_numpy(t).cuda(); do = torch.zeros(5 * n, dtype=torch.uint8, device='cuda')\n"
        "    for mb in (n // 3 + 1, 1000):\n"
        "        assert np.array_equal(c.construct_sa(t, max_block=mb), sa5), (name, mb)\n"
        "        do.zero_(); c.construct_sa_ptr(dt.data_ptr(), n, do.data_ptr(), mb, device=True)\n"
        "        assert np.array_equal(do.cpu().numpy(), sa5), (name, mb, 'device')\n"
        "print('paths ok', c.stats()['n_blocks'])\n" % (root, os.path.join(root, "tests")))
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, **env),
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "paths ok" in r.stdout, r.stderr[-2000:]


def test_work_guard_subprocess():
    """K0 and the initial-rank search cap every direct comparison and finish long ones by fingerprints, so a
    degenerate text (one repeated symbol) is sorted under a tiny comparison budget.  The symbol-by-symbol path
    with its device-side work guard remains for texts the fingerprints cannot cover -- a device text that is not
    16-byte aligned: there the same budget must end in PSC_ELIMIT (-5) instead of a multi-hour kernel, and the
    default budget must give the right answer."""
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys, numpy as np, torch; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "import psascan_b200 as P, oracle_lib as O\n"
        "from psascan_b200.lib import PscError\n"
        "n = 300000\n"
        "t = np.full(n, 65, dtype=np.uint8)\n"
        "want = O.encode_uint40(O.sa(t))\n"
        "c = P.Context(0)\n"
        "assert np.array_equal(c.construct_sa(t, max_block=100000), want)\n"
        "print('aligned ok')\n"
        "buf = torch.zeros(n + 64, dtype=torch.uint8, device='cuda'); buf[1:n + 1] = 65\n"
        "out = torch.zeros(5 * n, dtype=torch.uint8, device='cuda')\n"
        "try:\n"
        "    c.construct_sa_ptr(buf.data_ptr() + 1, n, out.data_ptr(), 100000, device=True)\n"
        "    assert np.array_equal(out.cpu().numpy(), want)\n"
        "    print('unaligned ok')\n"
        "except PscError as e:\n"
        "    print('code', e.code)\n" % (root, os.path.join(root, "tests")))
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, PSC_COMPARE_BUDGET="65536"),
                       capture_output=True, text=True, timeout=600)
    assert "aligned ok" in r.stdout and "code -5" in r.stdout, (r.stdout, r.stderr[-1500:])
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ), capture_output=True, text=True, timeout=600)
    assert "aligned ok" in r.stdout and "unaligned ok" in r.stdout, (r.stdout, r.stderr[-1500:])


def test_verify_sa_device(ctx):
    """psc_verify_sa_d accepts the suffix array and counts each kind of damage (out of range, duplicate,
    first-symbol order, tie order)."""
    import torch
    text = textgen.markov(30000, seed=21)
    n = len(text)
    sa5 = ctx.construct_sa(text, max_block=7000)
    d_text = torch.from_numpy(text).cuda()

    def check(buf):
        d = torch.from_numpy(buf).cuda()
        return ctx.verify_sa_ptr(d_text.data_ptr(), n, d.data_ptr())

    assert check(sa5) == [0, 0, 0, 0]
    sa = decode_sa5(sa5)
    bad = sa.copy(); bad[100] = n + 5
    v = check(O.encode_uint40(bad)); assert v[0] == 1
    bad = sa.copy(); bad[200] = bad[201]
    v = check(O.encode_uint40(bad)); assert v[1] == 1
    # swap two neighbours that start with the same symbol: only the tie order is wrong
    k = next(i for i in range(n - 1) if text[sa[i]] == text[sa[i + 1]])
    bad = sa.copy(); bad[k], bad[k + 1] = bad[k + 1], bad[k]
    v = check(O.encode_uint40(bad)); assert v[:2] == [0, 0] and v[3] >= 1
    k = next(i for i in range(n - 1) if text[sa[i]] != text[sa[i + 1]])
    bad = sa.copy(); bad[k], bad[k + 1] = bad[k + 1], bad[k]
    v = check(O.encode_uint40(bad)); assert v[:2] == [0, 0] and v[2] == 1


def test_multi_gpu_dist_check_torchrun():
    """csrc/dist.cu over NCCL on every visible GPU (at least 2): tools/dist_check.py compares both distributed
    entry points with the golden .sa5 files and the single-GPU pipeline, bit for bit, and exits non-zero on
    any difference."""
    import os, subprocess, sys, torch
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs at least 2 GPUs (one rank per GPU)")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    world = min(ngpu, 8)
    port = 29600 + os.getpid() % 300
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(root, "tools", "dist_check.py")],
                       capture_output=True, text=True, timeout=1200)
    assert r.returncode == 0, (r.stdout[-3000:], r.stderr[-3000:])
    assert "ok=False" not in r.stdout and r.stdout.count("ok=True") >= 10


def test_config1_full_size_cmp_with_reference(tmp_path_factory):
    """BASELINE configs[0] at its stated size: 200 MB of DNA-like text, host/construct_sa (GPU) against the
    unmodified reference oracle/_ref/construct_sa -m 3584mi (the "3.5gi" of BASELINE.json: the option takes whole numbers only) -- the two .sa5 files (10^9 bytes) must be equal."""
    import filecmp, os, shutil, subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ours, ref = os.path.join(root, "host", "construct_sa"), O.REF_BIN
    if not (os.path.exists(ours) and os.path.exists(ref)):
        pytest.skip("host/construct_sa or oracle/_ref/construct_sa not built")
    base = "/dev/shm" if os.path.isdir("/dev/shm") and shutil.disk_usage("/dev/shm").free > (3 << 30) else None
    d = tmp_path_factory.mktemp("cfg1") if base is None else None
    work = str(d) if d is not None else os.path.join(base, "psc_cfg1_%d" % os.getpid())
    os.makedirs(work, exist_ok=True)
    try:
        f = os.path.join(work, "dna200m.txt")
        textgen.dna(200_000_000, seed=1).tofile(f)
        threads = str(min(16, len(os.sched_getaffinity(0))))
        subprocess.run([ref, f, "-o", f + ".ref.sa5", "-m", "3584mi"], check=True, stdin=subprocess.DEVNULL,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, env=dict(os.environ, OMP_NUM_THREADS=threads),
                       timeout=1500)
        subprocess.run([ours, f, "-o", f + ".gpu.sa5", "-m", "3584mi"], check=True, stdin=subprocess.DEVNULL,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=600)
        assert os.path.getsize(f + ".gpu.sa5") == 1_000_000_000
        assert filecmp.cmp(f + ".ref.sa5", f + ".gpu.sa5", shallow=False)
    finally:
        shutil.rmtree(work, ignore_errors=True)


def test_config2_full_size_verified(ctx):
    """BASELINE configs[1] at its stated size: 2^30 uniform random bytes (0..254), one block; the result goes
    through the O(n) device check (no CPU sorter reaches this size inside a test)."""
    import torch
    n = 1 << 30
    g = torch.Generator(device="cuda"); g.manual_seed(2)
    text = torch.randint(0, 255, (n,), dtype=torch.uint8, device="cuda", generator=g)
    out = torch.empty(5 * n, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()   # the library works on its own non-blocking stream: inputs must be complete
    ctx.construct_sa_ptr(text.data_ptr(), n, out.data_ptr(), 0, device=True)
    assert ctx.stats()["n_blocks"] == 1
    assert ctx.verify_sa_ptr(text.data_ptr(), n, out.data_ptr()) == [0, 0, 0, 0]
    # and the same text in 5 ragged blocks: rank planes of 255 symbols, gap arrays, the merge by output range
    ctx.construct_sa_ptr(text.data_ptr(), 1 << 28, out.data_ptr(), (1 << 26) - 12345, device=True)
    assert ctx.stats()["n_blocks"] == 5
    assert ctx.verify_sa_ptr(text.data_ptr(), 1 << 28, out.data_ptr()) == [0, 0, 0, 0]
    # 3 ragged blocks of ~128 MiB: the BWT in passes over 64 MiB text windows (sufsort.cu, 2 full + 1 ragged)
    ctx.construct_sa_ptr(text.data_ptr(), 1 << 28, out.data_ptr(), (1 << 27) - 12345, device=True)
    assert ctx.stats()["n_blocks"] == 3
    assert ctx.verify_sa_ptr(text.data_ptr(), 1 << 28, out.data_ptr()) == [0, 0, 0, 0]
    del text, out
    torch.cuda.empty_cache()


def _run_child(code, env, timeout=1200):
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pre = "import sys, numpy as np, torch; sys.path.insert(0, %r); sys.path.insert(0, %r)\n" % (root, os.path.join(root, "tests"))
    return subprocess.run([sys.executable, "-c", pre + code], env=dict(os.environ, **env), capture_output=True, text=True,
                          timeout=timeout)


@pytest.mark.parametrize("cap", ["4", "64"])
def test_fingerprint_lcp_forced_on_golden_cases(cap):
    """PSC_LCP_CAP (read once per process) cut to a few symbols: nearly every K0 / K5 comparison that matters is
    finished by the Karp-Rabin fingerprints of texthash.cuh -- stage calls and whole constructions must still be
    bit-exact against the reference's golden files and the oracle."""
    code = (
        "import psascan_b200 as P, oracle_lib as O\n"
        "from golden_cases import golden_cases, golden_text\n"
        "from psascan_b200 import textgen\n"
        "c = P.Context(0)\n"
        "for name, gen, kw, n, sa5 in golden_cases():\n"
        "    t = golden_text(gen, kw, n)\n"
        "    for mb in (n // 2 + 1, n // 3 + 1, 1000):\n"
        "        assert np.array_equal(c.construct_sa(t, max_block=mb), sa5), (name, mb)\n"
        "        dt = torch.from_numpy(t).cuda(); do = torch.zeros(5 * n, dtype=torch.uint8, device='cuda')\n"
        "        c.construct_sa_dist_ptr(dt.data_ptr(), n, do.data_ptr(), mb)\n"
        "        assert np.array_equal(do.cpu().numpy(), sa5), (name, mb, 'dist: K0 past the block end')\n"
        "for t in (textgen.periodic(30000, period=7, sigma=2), np.zeros(20000, dtype=np.uint8), textgen.periodic(50000, period=997, sigma=3)):\n"
        "    want = O.encode_uint40(O.sa(t))\n"
        "    for mb in (len(t) // 4 + 1, 5000):\n"
        "        assert np.array_equal(c.construct_sa(t, max_block=mb), want), (len(t), mb)\n"
        "print('fingerprints ok')\n")
    r = _run_child(code, dict(PSC_LCP_CAP=cap, PSC_HASH_SEED="12345"))
    assert r.returncode == 0 and "fingerprints ok" in r.stdout, (r.stdout[-500:], r.stderr[-2500:])


def test_periodic_texts_across_block_boundaries_bounded():
    """Texts whose repeats across a block boundary are as long as the block (period 997; one repeated byte), in 4
    blocks, under a comparison budget that the symbol-by-symbol K0 / K5 of round 1 exceeded by orders of magnitude
    (PSC_COMPARE_BUDGET = 2^36).  Small sizes are compared with the unmodified reference binary, the sizes of
    VERDICT r1 item 8 (256 MiB / 64 MiB) go through the O(n) device verifier."""
    code = (
        "import psascan_b200 as P, oracle_lib as O\n"
        "from psascan_b200 import textgen\n"
        "c = P.Context(0)\n"
        "for t in (textgen.periodic(3 << 20, period=997, sigma=3), np.full(1 << 20, 97, dtype=np.uint8)):\n"
        "    got = c.construct_sa(t, max_block=len(t) // 4)\n"
        "    assert c.stats()['n_blocks'] == 4\n"
        "    assert np.array_equal(got, O.run_reference(t, threads=8)), len(t)\n"
        "for n, make in ((256 << 20, lambda n: torch.from_numpy(textgen.periodic(997, period=997, sigma=3)).cuda().repeat(n // 997 + 1)[:n].contiguous()),\n"
        "                (64 << 20, lambda n: torch.full((n,), 97, dtype=torch.uint8, device='cuda'))):\n"
        "    t = make(n); out = torch.empty(5 * n, dtype=torch.uint8, device='cuda'); torch.cuda.synchronize()\n"
        "    c.construct_sa_ptr(t.data_ptr(), n, out.data_ptr(), n // 4, device=True)\n"
        "    st = c.stats(); assert st['n_blocks'] == 4\n"
        "    assert c.verify_sa_ptr(t.data_ptr(), n, out.data_ptr()) == [0, 0, 0, 0], n\n"
        "    print('big', n, round(st['ms_total']), 'ms; gt_end', round(st['ms_gt_end']), 'init_ranks', round(st['ms_init_ranks']), 'sort', round(st['ms_sort']))\n"
        "    del t, out\n"
        "print('periodic ok')\n")
    r = _run_child(code, dict(PSC_COMPARE_BUDGET=str(1 << 36)), timeout=3000)
    assert r.returncode == 0 and "periodic ok" in r.stdout, (r.stdout[-800:], r.stderr[-2500:])


def test_verify_sa_slices(ctx):
    """psc_verify_sa_slice_d: slices of a correct suffix array are clean (also on a periodic text whose adjacent
    suffixes share prefixes far beyond the direct-compare cap), every kind of damage is counted, and the slice
    boundary pair is checked through next_first."""
    import torch
    for text in (textgen.markov(60000, seed=41), textgen.periodic(40000, period=13, sigma=2), np.zeros(5000, dtype=np.uint8)):
        n = len(text)
        sa5 = ctx.construct_sa(text, max_block=n // 3 + 1)
        sa = decode_sa5(sa5)
        d_text = torch.from_numpy(text).cuda()

        def check(arr, beg, ln):
            d = torch.from_numpy(O.encode_uint40(arr[beg:beg + ln])).cuda()
            nxt = int(arr[beg + ln]) if beg + ln < n else -1
            return ctx.verify_sa_slice_ptr(d_text.data_ptr(), n, d.data_ptr(), beg, ln, nxt)

        cuts = [0, n // 3, n // 3 + 1, 2 * n // 3, n]
        for a, b in zip(cuts, cuts[1:]):
            assert check(sa, a, b - a) == [0, 0], (n, a, b)
        bad = sa.copy(); bad[7] = n + 1
        assert check(bad, 0, n // 3)[0] == 1
        bad = sa.copy(); bad[100], bad[101] = bad[101], bad[100]
        assert check(bad, 0, n // 3)[1] >= 1
        bad = sa.copy(); bad[500] = bad[499]
        assert check(bad, 0, n)[1] >= 1
        k = n // 3   # damage exactly at a slice boundary: only next_first can see it
        bad = sa.copy(); bad[k - 1], bad[k] = bad[k], bad[k - 1]
        assert check(bad, 0, k)[1] + check(bad, k, n - k)[1] >= 1
