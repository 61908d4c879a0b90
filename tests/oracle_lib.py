"""ctypes binding of oracle/liboracle.so -- the CPU restatement (TEST INFRASTRUCTURE ONLY).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this."""
import ctypes as C
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_lib = None


def lib():
    global _lib
    if _lib is None:
        so = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(so):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "oracle"], stdout=subprocess.DEVNULL)
        _lib = C.CDLL(so)
        for f in ("orc_max_block_size", "orc_sort_block", "orc_initial_rank", "orc_compute_gap", "orc_merge_bwt"):
            getattr(_lib, f).restype = C.c_long
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def u8(a):
    return np.ascontiguousarray(a, dtype=np.uint8)


def sa_naive(text):
    text = u8(text)
    sa = np.zeros(len(text), dtype=np.int64)
    lib().orc_sa_naive(_p(text), C.c_long(len(text)), _p(sa))
    return sa


def sa(text):
    text = u8(text)
    out = np.zeros(len(text), dtype=np.int64)
    lib().orc_sa(_p(text), C.c_long(len(text)), _p(out))
    return out


def psascan(text, max_block, ram_use=1 << 40, half_blocks=0):
    text = u8(text)
    out = np.zeros(len(text), dtype=np.int64)
    rc = lib().orc_psascan(_p(text), C.c_long(len(text)), C.c_long(max_block), C.c_long(ram_use),
                           C.c_int(half_blocks), _p(out))
    return rc, out


def encode_uint40(sa_):
    sa_ = np.ascontiguousarray(sa_, dtype=np.int64)
    out = np.zeros(5 * len(sa_), dtype=np.uint8)
    lib().orc_encode_uint40(_p(sa_), C.c_long(len(sa_)), _p(out))
    return out


def decode_uint40(buf):
    buf = u8(buf)
    n = len(buf) // 5
    out = np.zeros(n, dtype=np.int64)
    lib().orc_decode_uint40(_p(buf), C.c_long(n), _p(out))
    return out


def gt_end(text, beg, end, gt_tail):
    text = u8(text)
    n = len(text)
    out = np.zeros(n + 2, dtype=np.uint8)
    lib().orc_gt_end(_p(text), C.c_long(n), C.c_long(beg), C.c_long(end), _p(u8(gt_tail)), _p(out))
    return out


def sort_block(text, beg, end, gt_tail):
    """returns psa_rel, bwt, gt_begin (flags by global position, len n+2), i0"""
    text = u8(text)
    n, m = len(text), end - beg
    psa = np.zeros(m, dtype=np.int64)
    bwt = np.zeros(m, dtype=np.uint8)
    gtb = np.zeros(n + 2, dtype=np.uint8)
    gt_tail = u8(gt_tail) if gt_tail is not None else np.zeros(n + 2, dtype=np.uint8)
    i0 = lib().orc_sort_block(_p(text), C.c_long(n), C.c_long(beg), C.c_long(end), _p(gt_tail),
                              _p(psa), _p(bwt), _p(gtb))
    return psa, bwt, gtb, i0


def initial_rank(text, beg, end, psa_rel, pat_beg, cmp_end, gt_cmp):
    text = u8(text)
    psa_rel = np.ascontiguousarray(psa_rel, dtype=np.int64)
    return lib().orc_initial_rank(_p(text), C.c_long(len(text)), C.c_long(beg), C.c_long(end), _p(psa_rel),
                                  C.c_long(pat_beg), C.c_long(cmp_end), _p(u8(gt_cmp)))


def compute_gap(text, bwt, i0, last, tail_beg, tail_end, start_rank, gt_in, want_gt_out=True):
    text, bwt = u8(text), u8(bwt)
    n, m = len(text), len(bwt)
    gap = np.zeros(m + 1, dtype=np.int64)
    gt_out = np.zeros(n + 2, dtype=np.uint8) if want_gt_out else None
    gt_in = u8(gt_in) if gt_in is not None else None
    r = lib().orc_compute_gap(_p(text), _p(bwt), C.c_long(m), C.c_long(i0), C.c_uint8(int(last)),
                              C.c_long(tail_beg), C.c_long(tail_end), C.c_long(start_rank), _p(gt_in),
                              _p(gt_out), _p(gap))
    return gap, gt_out, r


def gap_to_bv(gap_l, len_l, len_r):
    g = np.ascontiguousarray(gap_l, dtype=np.int64)
    bv = np.zeros(len_l + len_r + 1, dtype=np.uint8)
    lib().orc_gap_to_bv(_p(g), C.c_long(len_l), C.c_long(len_r), _p(bv))
    return bv[:len_l + len_r]


def merge_bwt(bwt_l, bwt_r, i0_l, i0_r, last_l, bv):
    l, r, bv = u8(bwt_l), u8(bwt_r), u8(bv)
    out = np.zeros(len(l) + len(r), dtype=np.uint8)
    i0 = lib().orc_merge_bwt(_p(l), _p(r), C.c_long(len(l)), C.c_long(len(r)), C.c_long(i0_l), C.c_long(i0_r),
                             C.c_uint8(int(last_l)), _p(bv), _p(out))
    return out, i0


def split_gap(bv, len_l, len_r, block_gap):
    bv = u8(bv)
    g = np.ascontiguousarray(block_gap, dtype=np.int64)
    lg = np.zeros(len_l + 1, dtype=np.int64)
    rg = np.zeros(len_r + 1, dtype=np.int64)
    lib().orc_split_gap(_p(bv), C.c_long(len_l), C.c_long(len_r), _p(g), _p(lg), _p(rg))
    return lg, rg


def merge(beg, length, psa_list, gap_list):
    k = len(beg)
    n = int(sum(length))
    b = np.ascontiguousarray(beg, dtype=np.int64)
    l = np.ascontiguousarray(length, dtype=np.int64)
    psa = [np.ascontiguousarray(x, dtype=np.int64) for x in psa_list]
    gap = [np.ascontiguousarray(x, dtype=np.int64) for x in gap_list]
    pp = (C.c_void_p * k)(*[x.ctypes.data for x in psa])
    pg = (C.c_void_p * k)(*[x.ctypes.data for x in gap])
    out = np.zeros(n, dtype=np.int64)
    lib().orc_merge(C.c_long(k), _p(b), _p(l), pp, pg, _p(out))
    return out


def pack_gt_rev(gt, n, lo, hi):
    gt = u8(gt)
    bits = np.zeros((hi - lo + 7) // 8 + 1, dtype=np.uint8)
    lib().orc_pack_gt_rev(_p(gt), C.c_long(n), C.c_long(lo), C.c_long(hi), _p(bits))
    return bits[:(hi - lo + 7) // 8]


def parse_number(s):
    v = C.c_uint64(0)
    ok = lib().orc_parse_number(s.encode(), C.byref(v))
    return (v.value if ok else None)


def max_block_size(ram, threads):
    return lib().orc_max_block_size(C.c_long(ram), C.c_long(threads))


REF_BIN = os.path.join(ORACLE_DIR, "_ref", "construct_sa")


def run_reference(text, mem=None, threads=2, workdir=None):
    """Runs the unmodified reference binary (oracle/_ref/construct_sa) on `text`; returns .sa5 bytes."""
    import tempfile
    d = workdir or tempfile.mkdtemp(prefix="psc_ref_")
    f = os.path.join(d, "t.txt")
    u8(text).tofile(f)
    out = f + ".sa5"
    if os.path.exists(out):
        os.remove(out)
    cmd = [REF_BIN, f, "-o", out]
    if mem:
        cmd += ["-m", str(mem)]
    env = dict(os.environ, OMP_NUM_THREADS=str(threads))
    subprocess.run(cmd, check=True, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL,
                   stdin=subprocess.DEVNULL)
    res = np.fromfile(out, dtype=np.uint8)
    for x in os.listdir(d):
        os.remove(os.path.join(d, x))
    if workdir is None:
        os.rmdir(d)
    return res
