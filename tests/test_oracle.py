"""CPU tests: pin the oracle (oracle/psascan_oracle.c) against the reference's outputs
(tests/golden/*.npz, written by the unmodified reference binary) and against brute force."""
import numpy as np
import pytest
import oracle_lib as O
from golden_cases import golden_cases, golden_text
from psascan_b200 import textgen

CASES = golden_cases()


def test_golden_fixtures_present():
    assert len(CASES) >= 8


@pytest.mark.parametrize("name,gen,kw,n,sa5", CASES, ids=[c[0] for c in CASES])
def test_oracle_matches_reference_sa5(name, gen, kw, n, sa5):
    text = golden_text(gen, kw, n)
    ref = O.decode_uint40(sa5)
    assert np.array_equal(O.sa(text), ref)
    assert np.array_equal(O.encode_uint40(ref), sa5)
    # block pipeline restatement: collapsed schedule (what the GPU runs) and the reference's
    # half-block schedule, several block sizes
    for mb in (n, n // 2 + 1, n // 3 + 7, 977):
        rc, sa = O.psascan(text, mb, half_blocks=0)
        assert rc == 0 and np.array_equal(sa, ref), (name, mb)
    rc, sa = O.psascan(text, n // 3 + 7, ram_use=10 * 1500, half_blocks=1)
    assert rc == 0 and np.array_equal(sa, ref)


@pytest.mark.parametrize("seed", range(6))
def test_oracle_vs_bruteforce(seed):
    rng = np.random.default_rng(seed)
    n = int(rng.integers(1, 400))
    sigma = int(rng.choice([1, 2, 3, 4, 26, 255]))
    text = rng.integers(0, sigma, size=n, dtype=np.uint8)
    naive = O.sa_naive(text)
    assert np.array_equal(O.sa(text), naive)
    for mb in (n, max(1, n // 2), max(1, n // 5), 1, 7):
        for hb in (0, 1):
            rc, sa = O.psascan(text, mb, ram_use=10 * max(1, n // 7), half_blocks=hb)
            assert rc == 0 and np.array_equal(sa, naive), (n, sigma, mb, hb)


def test_sentinel_rejected():
    text = np.array([1, 2, 255, 3, 4, 5], dtype=np.uint8)
    rc, _ = O.psascan(text, 3)
    assert rc == -1


def test_parse_number_and_budget():
    # src/main.cpp:76-131
    assert O.parse_number("3584Mi") == 3584 << 20
    assert O.parse_number("3.5gi") is None
    assert O.parse_number("10k") == 10000
    assert O.parse_number("7G") == 7 * 10**9
    assert O.parse_number("2ti") == 2 << 40
    assert O.parse_number("12x") is None and O.parse_number("mi") is None and O.parse_number("1kib") is None
    # include/psascan.hpp:73-91, values seen in reference logs: 3584 MiB, 8 threads
    mbs = O.max_block_size(3584 << 20, 8)
    assert abs(mbs / (1 << 20) - 667.7) < 0.1
    assert O.max_block_size(3584 << 20, 256) == -2


def test_uint40_roundtrip():
    v = np.array([0, 1, 255, 256, 2**32 - 1, 2**32, 2**40 - 1], dtype=np.int64)
    b = O.encode_uint40(v)
    assert len(b) == 35 and list(b[20:25]) == [255, 255, 255, 255, 0] and list(b[25:30]) == [0, 0, 0, 0, 1]
    assert np.array_equal(O.decode_uint40(b), v)


def test_stage_functions_compose():
    """stage-level oracle functions reproduce the pipeline by hand (2 blocks)."""
    text = textgen.periodic(700, period=31, sigma=3, seed=5)
    n, mid = len(text), 350
    ref = O.sa_naive(text)
    z = np.zeros(n + 2, dtype=np.uint8)
    psa1, bwt1, gtb1, i01 = O.sort_block(text, mid, n, z)
    psa0, bwt0, gtb0, i00 = O.sort_block(text, 0, mid, gtb1)
    gap0, gt_out, r = O.compute_gap(text, bwt0, i00, text[mid - 1], mid, n, 0, gtb1)
    assert gap0.sum() == n - mid
    # initial rank of an interior tail position equals the walk's rank there
    for p in (mid, mid + 64, n - 1):
        g2, _, rp = O.compute_gap(text, bwt0, i00, text[mid - 1], p, n, 0, gtb1)
        assert O.initial_rank(text, 0, mid, psa0, p, mid, gtb1) == rp
    sa = O.merge([0, mid], [mid, n - mid], [psa0, psa1], [gap0, np.zeros(n - mid + 1, dtype=np.int64)])
    assert np.array_equal(sa, ref)
    # gt_out: suffix(j) > suffix(0)
    isa = np.empty(n, dtype=np.int64); isa[ref] = np.arange(n)
    for j in range(mid + 1, n):
        assert gt_out[j] == (isa[j] > isa[0])
