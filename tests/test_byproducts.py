"""By-products through the C ABI: the sentinel filter against numpy and against the reference's
own tool (oracle/_ref/delete_sentinel_bytes, built from /root/reference by oracle/Makefile), the full-text BWT
and ISA samples against a brute-force restatement on the oracle's suffix array."""
import ctypes as C
import os
import subprocess
import numpy as np
import pytest
import oracle_lib as O
from psascan_b200 import textgen

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_delete_sentinel_bytes_device(ctx):
    import torch
    rng = np.random.default_rng(5)
    for n, p255 in ((1, 0.0), (1, 1.0), (4095, 0.3), (4096, 0.5), (100_001, 0.01), (3_000_000, 0.9), (1 << 20, 0.0)):
        a = rng.integers(0, 255, size=n, dtype=np.uint8)
        a[rng.random(n) < p255] = 255
        d_in = torch.from_numpy(a).cuda()
        d_out = torch.zeros(n, dtype=torch.uint8, device="cuda")
        kept = C.c_long(-1)
        ctx._chk(ctx.lib.psc_delete_sentinel_bytes_d(ctx.h, C.c_void_p(d_in.data_ptr()), C.c_long(n),
                                                     C.c_void_p(d_out.data_ptr()), C.byref(kept)))
        want = a[a != 255]
        assert kept.value == len(want)
        assert np.array_equal(d_out.cpu().numpy()[:kept.value], want), (n, p255)


def test_delete_sentinel_bytes_tool_vs_reference_tool(tmp_path):
    ours = os.path.join(ROOT, "host", "delete_sentinel_bytes")
    ref = os.path.join(ROOT, "oracle", "_ref", "delete_sentinel_bytes")
    if not (os.path.exists(ours) and os.path.exists(ref)):
        pytest.skip("tools not built")
    rng = np.random.default_rng(6)
    a = rng.integers(0, 256, size=150_000_000, dtype=np.uint8)   # several 64 MiB chunks, 255s included
    f = tmp_path / "in.bin"
    a.tofile(str(f))
    got = subprocess.run([ours, str(f)], capture_output=True, timeout=300)
    want = subprocess.run([ref, str(f)], capture_output=True, timeout=300)
    assert got.returncode == 0 and want.returncode == 0
    assert got.stdout == want.stdout and len(got.stdout) == int((a != 255).sum())
    assert b"Processed" in got.stderr and b"100.0%" in got.stderr
    r = subprocess.run([ours], capture_output=True)
    r2 = subprocess.run([ref], capture_output=True)
    assert r.returncode == r2.returncode == 1 and b"Erase all bytes with value 255" in r.stderr


def test_full_text_bwt_and_isa_samples(ctx):
    import torch
    for text in (textgen.markov(50_000, seed=3), textgen.dna(20_001, seed=4), np.array([7], dtype=np.uint8)):
        n = len(text)
        sa = O.sa(text)
        sa5 = ctx.construct_sa(text, max_block=max(1, n // 3 + 1))
        d_text, d_sa5 = torch.from_numpy(text).cuda(), torch.from_numpy(sa5).cuda()
        d_bwt = torch.zeros(n, dtype=torch.uint8, device="cuda")
        i0 = C.c_long(-1)
        ctx._chk(ctx.lib.psc_bwt_from_sa5_d(ctx.h, C.c_void_p(d_text.data_ptr()), C.c_long(n), C.c_void_p(d_sa5.data_ptr()),
                                            C.c_void_p(d_bwt.data_ptr()), C.byref(i0)))
        want = np.where(sa > 0, text[np.maximum(sa, 1) - 1], 0).astype(np.uint8)
        assert np.array_equal(d_bwt.cpu().numpy(), want)
        assert i0.value == int(np.nonzero(sa == 0)[0][0])
        for rate in (1, 7, 1024):
            ns = (n - 1) // rate + 1
            d_s = torch.zeros(ns, dtype=torch.int64, device="cuda")
            ctx._chk(ctx.lib.psc_isa_samples_d(ctx.h, C.c_void_p(d_sa5.data_ptr()), C.c_long(n), C.c_long(rate),
                                               C.c_void_p(d_s.data_ptr())))
            isa = np.empty(n, dtype=np.int64); isa[sa] = np.arange(n)
            assert np.array_equal(d_s.cpu().numpy(), isa[::rate])
