"""psc_verify_sa_noalloc_d (the allocation-free variant of the device-side SA check, static scratch inside the
library) against psc_verify_sa_d, through the C ABI, on two calls in a row.  Integer counters: exact.
Collected last on purpose (a failure here hides nothing behind it)."""
import ctypes as C
import numpy as np
import pytest
import oracle_lib as O
from psascan_b200 import textgen
from psascan_b200.lib import decode_sa5, PscError

pytestmark = pytest.mark.gpu


def test_verify_noalloc_matches_verify(ctx):
    import torch
    text = textgen.markov(150_000, seed=31)
    n = len(text)
    sa5 = ctx.construct_sa(text, max_block=40_000)
    sa = decode_sa5(sa5)
    bad = sa.copy()
    bad[10], bad[11] = bad[11], bad[10]      # order damage
    bad[500] = bad[501]                      # a duplicate (and a missing value)
    bad[9000] = n + 3                        # out of range
    small = textgen.dna(5_000, seed=32)      # a shorter input after a longer one: stale scratch must not leak in
    small_sa5 = ctx.construct_sa(small)
    d_text = torch.from_numpy(text).cuda()
    d_small = torch.from_numpy(small).cuda()
    cases = [(d_text, n, torch.from_numpy(sa5).cuda()),
             (d_text, n, torch.from_numpy(O.encode_uint40(bad)).cuda()),
             (d_small, len(small), torch.from_numpy(small_sa5).cuda()),
             (d_text, n, torch.from_numpy(sa5).cuda())]
    for rep in range(2):                     # two calls in a row per case, and the whole list twice
        for t, m, s in cases:
            want = ctx.verify_sa_ptr(t.data_ptr(), m, s.data_ptr())
            got1 = ctx.verify_sa_noalloc_ptr(t.data_ptr(), m, s.data_ptr())
            got2 = ctx.verify_sa_noalloc_ptr(t.data_ptr(), m, s.data_ptr())
            assert got1 == want and got2 == want, (rep, m, want, got1, got2)
    assert ctx.verify_sa_noalloc_ptr(d_text.data_ptr(), n, cases[0][2].data_ptr()) == [0, 0, 0, 0]
    assert any(ctx.verify_sa_noalloc_ptr(d_text.data_ptr(), n, cases[1][2].data_ptr()))


def test_verify_noalloc_rejects_sizes_beyond_the_scratch(ctx):
    ctx.lib.psc_verify_noalloc_max_n.restype = C.c_long
    limit = int(ctx.lib.psc_verify_noalloc_max_n())
    assert limit == 4194304
    with pytest.raises(PscError) as e:       # rejected before any pointer is touched
        ctx.verify_sa_noalloc_ptr(8, limit + 1, 8)
    assert e.value.code == -1 and "exceeds the static scratch" in str(e.value)
    assert ctx.verify_sa_noalloc_ptr(0, 0, 0) == [0, 0, 0, 0]
