"""ctypes binding of libpsascan_cuda.so (include/psascan_cuda.h).  Mirrors the reference seams:
sort_block ~ inmem_psascan(), Rank ~ rank4n, initial_ranks ~ em_compute_initial_ranks(),
compute_gap ~ compute_gap(), merge_level ~ merge(), construct_sa ~ pSAscan() in memory."""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
EXPORTS = [
    "psc_last_error", "psc_version", "psc_device_count", "psc_init", "psc_shutdown", "psc_max_block_size",
    "psc_sort_block", "psc_rank_build", "psc_rank_counts", "psc_rank_query", "psc_rank_free",
    "psc_initial_ranks", "psc_compute_gap", "psc_merge_level", "psc_gap_to_bitvector",
    "psc_merge_bwt", "psc_split_gap", "psc_construct_sa", "psc_construct_sa_d", "psc_get_stats",
    "psc_bench_random_sectors", "psc_bench_radix_pass", "psc_device_alloc", "psc_device_free",
    "psc_pinned_alloc", "psc_pinned_free", "psc_copy_h2d", "psc_copy_d2h", "psc_device_info", "psc_construct_sa_dist_d",
    "psc_construct_sa_dist_slice_d", "psc_dist_slice_capacity", "psc_verify_sa_d",
    "psc_verify_sa_noalloc_d", "psc_verify_noalloc_max_n", "psc_dist_unique_id", "psc_dist_init",
    "psc_dist_finalize", "psc_dist_stats", "psc_dist_abort", "psc_delete_sentinel_bytes_d",
    "psc_bwt_from_sa5_d", "psc_isa_samples_d", "psc_verify_sa_slice_d", "psc_construct_sa_files_d",
    "psc_max_block_size_files", "psc_vbyte_encode", "psc_vbyte_decode", "psc_multifile_read_bits",
]


class PscError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"psascan_cuda error {code}: {msg}")
        self.code = code


class Stats(C.Structure):
    _fields_ = [(k, C.c_double) for k in (
        "ms_total", "ms_gt_end", "ms_sort", "ms_radix", "ms_bwt", "ms_rank_build", "ms_init_ranks",
        "ms_walk", "ms_merge", "ms_h2d", "ms_d2h", "ms_radix_pass")] + [(k, C.c_long) for k in (
        "n_blocks", "lf_steps", "radix_passes", "radix_elems", "doubling_rounds", "walkers",
        "kernel_launches", "radix_bytes", "radix_keygen_bytes")] + [("ms_radix_keygen", C.c_double), ("ms_walk_kernel", C.c_double)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


def library_path():
    return os.path.join(_HERE, "libpsascan_cuda.so")


_lib = None


def load_library():
    """Loads the CUDA library; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        p = library_path()
        if not os.path.exists(p):
            raise ImportError(f"{p} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`"
                              " (there is no CPU fallback)")
        lib = C.CDLL(p)
        lib.psc_last_error.restype = C.c_char_p
        lib.psc_max_block_size.restype = C.c_long
        lib.psc_max_block_size.argtypes = [C.c_void_p, C.c_long]
        _lib = lib
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _u8(a):
    return np.ascontiguousarray(a, dtype=np.uint8)


def _longs(a):
    return np.ascontiguousarray(a, dtype=np.int64)


class Rank:
    def __init__(self, ctx, handle, m):
        self.ctx, self.handle, self.m = ctx, handle, m

    def counts(self):
        out = np.zeros(256, dtype=np.int64)
        self.ctx._chk(self.ctx.lib.psc_rank_counts(self.handle, _p(out)))
        return out

    def query(self, i, c):
        i, c = _longs(i), _u8(c)
        out = np.zeros(len(i), dtype=np.int64)
        self.ctx._chk(self.ctx.lib.psc_rank_query(self.ctx.h, self.handle, C.c_long(len(i)), _p(i), _p(c), _p(out)))
        return out

    def free(self):
        if self.handle:
            self.ctx.lib.psc_rank_free(self.ctx.h, self.handle)
            self.handle = None


class Context:
    def __init__(self, device=-1, hbm_budget=0):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.psc_init(C.c_int(device), C.c_size_t(hbm_budget), C.byref(h))
        if rc != 0:
            raise PscError(rc, self.lib.psc_last_error().decode())
        self.h = h

    def _chk(self, rc):
        if rc != 0:
            raise PscError(rc, self.lib.psc_last_error().decode())

    def close(self):
        if getattr(self, "h", None):
            self.lib.psc_shutdown(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def max_block_size(self, n):
        return int(self.lib.psc_max_block_size(self.h, C.c_long(n)))

    def stats(self):
        s = Stats()
        self._chk(self.lib.psc_get_stats(self.h, C.byref(s)))
        return s.as_dict()

    # ---- stage 1
    def sort_block(self, text, tail_prefix=None, tail_gt_rev=None, want_bwt=True, want_gt=True):
        text = _u8(text)
        m = len(text)
        tp = _u8(tail_prefix) if tail_prefix is not None and len(tail_prefix) else None
        tg = _u8(tail_gt_rev) if tp is not None else None
        sa = np.zeros(m, dtype=np.uint32)
        bwt = np.zeros(m, dtype=np.uint8) if want_bwt else None
        gt = np.zeros((m + 7) // 8, dtype=np.uint8) if want_gt else None
        i0 = C.c_long(-1)
        self._chk(self.lib.psc_sort_block(self.h, _p(text), C.c_long(m), _p(tp),
                                          C.c_long(0 if tp is None else len(tp)), _p(tg), _p(sa),
                                          _p(bwt), _p(gt), C.byref(i0)))
        return sa, bwt, gt, i0.value

    # ---- stage 3a
    def rank_build(self, bwt, i0=-1):
        bwt = _u8(bwt)
        h = C.c_void_p()
        self._chk(self.lib.psc_rank_build(self.h, _p(bwt), C.c_long(len(bwt)), C.c_long(i0), C.byref(h)))
        return Rank(self, h, len(bwt))

    # ---- stage 2
    def initial_ranks(self, block_text, block_sa, after_text, after_gt_rev, starts):
        bt, at = _u8(block_text), _u8(after_text)
        sa = np.ascontiguousarray(block_sa, dtype=np.uint32)
        st = _longs(starts)
        g = _u8(after_gt_rev) if len(at) else None
        out = np.zeros(len(st), dtype=np.int64)
        self._chk(self.lib.psc_initial_ranks(self.h, _p(bt), _p(sa), C.c_long(len(bt)),
                                             _p(at) if len(at) else None, C.c_long(len(at)), _p(g),
                                             C.c_long(len(st)), _p(st), _p(out)))
        return out

    # ---- stage 3b
    def compute_gap(self, rank, i0, last, block_counts, tail, walk_beg, start_rank, gt_in_rev,
                    want_gt_out=True, gap=None):
        tail = _u8(tail)
        t = len(tail)
        wb, sr, bc = _longs(walk_beg), _longs(start_rank), _longs(block_counts)
        gin = _u8(gt_in_rev) if gt_in_rev is not None else None
        gout = np.zeros((t + 7) // 8, dtype=np.uint8) if want_gt_out else None
        if gap is None:
            gap = np.zeros(rank.m + 1, dtype=np.uint32)
        self._chk(self.lib.psc_compute_gap(self.h, rank.handle, C.c_long(i0), C.c_uint8(int(last)), _p(bc),
                                           _p(tail), C.c_long(t), C.c_long(len(wb) - 1), _p(wb), _p(sr),
                                           _p(gin), _p(gout), _p(gap)))
        return gap, gout

    # ---- stage 4
    def merge_level(self, sa, gap, beg, right_sa5):
        sa = np.ascontiguousarray(sa, dtype=np.uint32)
        m = len(sa)
        r5 = _u8(right_sa5) if right_sa5 is not None and len(right_sa5) else None
        rl = 0 if r5 is None else len(r5) // 5
        g = np.ascontiguousarray(gap, dtype=np.uint32) if rl else None
        out = np.zeros(5 * (m + rl), dtype=np.uint8)
        self._chk(self.lib.psc_merge_level(self.h, _p(sa), _p(g), C.c_long(m), C.c_long(beg), _p(r5),
                                           C.c_long(rl), _p(out)))
        return out

    # ---- glue
    def gap_to_bitvector(self, gap_l, len_l, len_r):
        g = np.ascontiguousarray(gap_l, dtype=np.uint32)
        bv = np.zeros((len_l + len_r + 7) // 8, dtype=np.uint8)
        self._chk(self.lib.psc_gap_to_bitvector(self.h, _p(g), C.c_long(len_l), C.c_long(len_r), _p(bv)))
        return bv

    def merge_bwt(self, bwt_l, bwt_r, i0_l, i0_r, last_l, bv):
        l, r, bv = _u8(bwt_l), _u8(bwt_r), _u8(bv)
        out = np.zeros(len(l) + len(r), dtype=np.uint8)
        i0 = C.c_long(-1)
        self._chk(self.lib.psc_merge_bwt(self.h, _p(l), _p(r), C.c_long(len(l)), C.c_long(len(r)),
                                         C.c_long(i0_l), C.c_long(i0_r), C.c_uint8(int(last_l)), _p(bv),
                                         _p(out), C.byref(i0)))
        return out, i0.value

    def split_gap(self, bv, len_l, len_r, block_gap):
        bv = _u8(bv)
        g = np.ascontiguousarray(block_gap, dtype=np.uint32)
        lg = np.zeros(len_l + 1, dtype=np.uint64)
        rg = np.zeros(len_r + 1, dtype=np.uint64)
        self._chk(self.lib.psc_split_gap(self.h, _p(bv), C.c_long(len_l), C.c_long(len_r), _p(g), _p(lg), _p(rg)))
        return lg, rg

    # ---- whole path
    def construct_sa(self, text, max_block=0, out=None):
        """text (numpy uint8, host) -> .sa5 bytes (numpy uint8, 5n).  Host buffers in and out."""
        text = _u8(text)
        n = len(text)
        if out is None:
            out = np.empty(5 * n, dtype=np.uint8)
        self._chk(self.lib.psc_construct_sa(self.h, _p(text), C.c_long(n), C.c_long(max_block), _p(out)))
        return out

    def construct_sa_ptr(self, text_ptr, n, out_ptr, max_block=0, device=False):
        """raw-pointer variant: device=True takes device pointers (HBM-resident timing)."""
        f = self.lib.psc_construct_sa_d if device else self.lib.psc_construct_sa
        self._chk(f(self.h, C.c_void_p(text_ptr), C.c_long(n), C.c_long(max_block), C.c_void_p(out_ptr)))

    def construct_sa_dist_ptr(self, text_dev_ptr, n, out_dev_ptr, max_block):
        """csrc/dist.cu schedule; a context that never joined a communicator (psascan_b200.dist.init) is a world of one"""
        self._chk(self.lib.psc_construct_sa_dist_d(self.h, C.c_void_p(text_dev_ptr), C.c_long(n), C.c_long(max_block),
                                                   C.c_void_p(out_dev_ptr)))

    def verify_sa_ptr(self, text_dev_ptr, n, sa5_dev_ptr):
        """O(n) device check (psc_verify_sa_d): returns the four violation counts, all zero for a correct SA"""
        v = (C.c_long * 4)()
        self._chk(self.lib.psc_verify_sa_d(self.h, C.c_void_p(text_dev_ptr), C.c_long(n), C.c_void_p(sa5_dev_ptr), v))
        return [int(x) for x in v]

    def verify_sa_slice_ptr(self, text_dev_ptr, n, slice_dev_ptr, slice_beg, slice_len, next_first=-1):
        """psc_verify_sa_slice_d: [entries >= n, adjacent pairs out of suffix order] for one slice of the SA"""
        v = (C.c_long * 2)()
        self._chk(self.lib.psc_verify_sa_slice_d(self.h, C.c_void_p(text_dev_ptr), C.c_long(n), C.c_void_p(slice_dev_ptr),
                                                 C.c_long(slice_beg), C.c_long(slice_len), C.c_long(next_first), v))
        return [int(x) for x in v]

    def verify_sa_noalloc_ptr(self, text_dev_ptr, n, sa5_dev_ptr):
        """psc_verify_sa_noalloc_d: the same check with the library's static scratch (no allocation, n bounded)"""
        v = (C.c_long * 4)()
        self._chk(self.lib.psc_verify_sa_noalloc_d(self.h, C.c_void_p(text_dev_ptr), C.c_long(n),
                                                   C.c_void_p(sa5_dev_ptr), v))
        return [int(x) for x in v]

    def bench_random_sectors(self, nbytes, iters=256):
        g = C.c_double(0)
        self._chk(self.lib.psc_bench_random_sectors(self.h, C.c_size_t(nbytes), C.c_int(iters), C.byref(g)))
        return g.value

    def bench_radix_pass(self, n, iters=8):
        g = C.c_double(0)
        self._chk(self.lib.psc_bench_radix_pass(self.h, C.c_long(n), C.c_int(iters), C.byref(g)))
        return g.value


def decode_sa5(buf):
    """uint40 little-endian entries -> int64 (types/uint40.hpp:42-46)."""
    b = np.asarray(buf, dtype=np.uint8).reshape(-1, 5).astype(np.int64)
    return b[:, 0] | (b[:, 1] << 8) | (b[:, 2] << 16) | (b[:, 3] << 24) | (b[:, 4] << 32)
