// em.cu -- external-memory mode: partial suffix arrays and gap arrays leave the device
// as FILES in the reference's own formats, and the final merge streams them back.
//
//   per block  SA of the block as block-relative int32 items in `OUT.distrfile.<hash>.part<k>` files of at most
//              part_bytes each (io/distributed_file.hpp:51-125, created at partial_sufsort.hpp:206-208, 331-333)
//              and its gap array v-byte coded in `GAP.gap.<hash>` (gap_array.hpp:156-182; 7 bits per byte, low
//              group first, high bit = "more"), encoded on the device;
//   merge      merge.hpp:54-180 reads every part and gap file front to back.  So does this merge: output chunk
//              after output chunk (finalmerge.cu), and because the slice a chunk needs from a block starts where
//              the previous chunk's slice ended, every file is read sequentially, once.  The split of the chunk
//              boundaries over the blocks only needs the coarse carries of every gap array (one sum per 2048
//              counters, kept in host memory with the byte offset of that tile in the gap file) plus one tile of
//              the file per boundary and block.  Files are deleted at the end like the reference does
//              (merge.hpp:178-179) unless PSC_KEEP_PARTS is set.
// Device memory: the text, one block's working set and one output chunk -- the 8 bytes per suffix that the
// resident pipeline keeps until the merge are on disk.
#include <errno.h>
#include <unistd.h>
#include <algorithm>
#include <random>
#include <string>
#include <vector>
#include "finalmerge.cuh"
#include "scan.cuh"

namespace psc {

// ---- v-byte on the device ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    vbyte_len_kernel(const uint32_t *__restrict__ v, long count, uint32_t *__restrict__ len) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const uint32_t x = v[i];
  len[i] = x < (1u << 7) ? 1u : x < (1u << 14) ? 2u : x < (1u << 21) ? 3u : x < (1u << 28) ? 4u : 5u;
}

__global__ void __launch_bounds__(256)
    vbyte_write_kernel(const uint32_t *__restrict__ v, const uint64_t *__restrict__ off, long count,
                       uint8_t *__restrict__ out) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  uint32_t x = v[i];
  uint8_t *o = out + off[i];
  while (x > 127) { *o++ = (uint8_t)((x & 0x7f) | 0x80); x >>= 7; }
  *o = (uint8_t)x;
}

// a value starts at byte j when j == 0 or byte j - 1 has no continuation bit
__global__ void __launch_bounds__(256)
    vbyte_mark_kernel(const uint8_t *__restrict__ b, long nbytes, uint32_t *__restrict__ starts) {
  const long j = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nbytes) return;
  starts[j] = (j == 0 || !(b[j - 1] & 0x80)) ? 1u : 0u;
}

// idx[j] = number of value starts before byte j.  complete[0] counts the values whose last byte is inside the
// buffer; pos_of_want = byte offset at which value number `want` starts.
__global__ void __launch_bounds__(256)
    vbyte_decode_kernel(const uint8_t *__restrict__ b, long nbytes, const uint32_t *__restrict__ starts,
                        const uint64_t *__restrict__ idx, long cap, uint32_t *__restrict__ out, long want,
                        unsigned long long *__restrict__ info /* [0] complete values, [1] byte of value `want` */) {
  const long j = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nbytes || !starts[j]) return;
  const long v = (long)idx[j];
  if (v == want) info[1] = (unsigned long long)j;
  uint64_t x = 0;
  int shift = 0;
  long k = j;
  while (k < nbytes && (b[k] & 0x80)) { x |= (uint64_t)(b[k] & 0x7f) << shift; shift += 7; ++k; }
  if (k >= nbytes) return;  // cut off by the end of the buffer
  x |= (uint64_t)b[k] << shift;
  atomicAdd(&info[0], 1ull);
  if (v < cap) out[v] = (uint32_t)x;
}

// gap_d[count] -> bytes (device); tile_off[t] = byte offset of entry t * COARSE_TILE, tile_off[last] = total
static size_t vbyte_encode_d(psc_context *ctx, const uint32_t *gap_d, long count, DevBuf<uint8_t> &bytes,
                             std::vector<uint64_t> *tile_off) {
  DevBuf<uint32_t> len(ctx, count);
  DevBuf<uint64_t> off(ctx, count + 1), total(ctx, 1);
  vbyte_len_kernel<<<(unsigned)div_up(count, 256), 256, 0, ctx->stream>>>(gap_d, count, len.p);
  PSC_LAUNCH_CHECK();
  exclusive_scan_u32_to_u64(ctx, len.p, off.p, count, total.p);
  unsigned long long nb = 0;
  d2h(ctx, &nb, reinterpret_cast<unsigned long long *>(total.p), 1);
  sync(ctx);
  bytes.alloc(ctx, (size_t)nb);
  vbyte_write_kernel<<<(unsigned)div_up(count, 256), 256, 0, ctx->stream>>>(gap_d, off.p, count, bytes.p);
  PSC_LAUNCH_CHECK();
  ctx->launches += 2;
  if (tile_off) {
    const long tiles = coarse_carry_tiles(count);
    tile_off->assign(tiles + 1, (uint64_t)nb);
    // every COARSE_TILE-th offset: strided copy
    PSC_CUDA(cudaMemcpy2DAsync(tile_off->data(), 8, off.p, (size_t)COARSE_TILE * 8, 8, (size_t)tiles, cudaMemcpyDeviceToHost, ctx->stream));
    sync(ctx);
    (*tile_off)[tiles] = (uint64_t)nb;
  }
  return (size_t)nb;
}

// decodes up to `cap` values of bytes_d[nbytes] into out_d; returns the number of complete values in the buffer and
// the byte offset at which value `want` starts (nbytes when want == number of complete values and none is cut off)
static long vbyte_decode_d(psc_context *ctx, const uint8_t *bytes_d, long nbytes, long cap, uint32_t *out_d, long want,
                           size_t *byte_of_want) {
  if (nbytes <= 0) { if (byte_of_want) *byte_of_want = 0; return 0; }
  DevBuf<uint32_t> starts(ctx, nbytes);
  DevBuf<uint64_t> idx(ctx, nbytes);
  DevBuf<unsigned long long> info(ctx, 2);
  unsigned long long init[2] = {0ull, (unsigned long long)nbytes};
  h2d(ctx, info.p, init, 2);
  vbyte_mark_kernel<<<(unsigned)div_up(nbytes, 256), 256, 0, ctx->stream>>>(bytes_d, nbytes, starts.p);
  PSC_LAUNCH_CHECK();
  exclusive_scan_u32_to_u64(ctx, starts.p, idx.p, nbytes, nullptr);
  vbyte_decode_kernel<<<(unsigned)div_up(nbytes, 256), 256, 0, ctx->stream>>>(bytes_d, nbytes, starts.p, idx.p, cap, out_d, want, info.p);
  PSC_LAUNCH_CHECK();
  ctx->launches += 2;
  unsigned long long h[2];
  d2h(ctx, h, info.p, 2);
  sync(ctx);
  if (byte_of_want) *byte_of_want = (size_t)h[1];
  return (long)h[0];
}

// ---- files ----------------------------------------------------------------------------------------------------
struct EmBlock {
  long beg = 0, m = 0;
  std::string sa_base;  // OUT.distrfile.<hash>; parts are sa_base + ".part" + k
  long items_per_part = 1;
  int n_parts = 0;
  std::string gap_file;         // empty for the rightmost block
  std::string gt_file;          // gt_begin bits of the block (PSC_KEEP_PARTS only)
  std::vector<uint64_t> carry;  // sum of the first t * COARSE_TILE counters, t = 0..tiles (gap_array.hpp sums nothing: ours)
  std::vector<uint64_t> voff;   // byte offset of counter t * COARSE_TILE in the gap file
  // cursors of the streaming merge
  long sa_next = 0;
  int sa_part = -1;
  FILE *sa_fp = nullptr, *gap_fp = nullptr;
  size_t gap_byte = 0;
};

static void io_fail(const char *what, const std::string &name) {
  set_error("%s %s: %s", what, name.c_str(), strerror(errno));
  throw CudaError{PSC_EARG};
}

static std::string random_hash() {  // utils.cpp:147-152 gives such names to temporary files
  static std::random_device rd;
  return std::to_string(((uint64_t)rd() << 31) ^ rd());
}

struct PinnedBuf {
  uint8_t *p = nullptr;
  size_t bytes = 0;
  explicit PinnedBuf(size_t b) : bytes(b) { PSC_CUDA(cudaMallocHost(reinterpret_cast<void **>(&p), b)); }
  ~PinnedBuf() { if (p) cudaFreeHost(p); }
};

struct EmSink : PieceSink {
  std::string out_path, gap_prefix;
  long n = 0, part_bytes = 0;
  std::vector<EmBlock> blocks;
  PinnedBuf stage;
  double ms_write = 0;
  EmSink(const std::string &out, const std::string &gap, long n_, long B, long part_bytes_)
      : out_path(out), gap_prefix(gap), n(n_), part_bytes(part_bytes_), blocks(B), stage((size_t)64 << 20) {}

  // device bytes -> file, through the pinned buffer
  void write_from_device(psc_context *ctx, const void *src_d, size_t bytes, FILE *f, const std::string &name) {
    for (size_t done = 0; done < bytes;) {
      const size_t now = std::min(stage.bytes, bytes - done);
      PSC_CUDA(cudaMemcpyAsync(stage.p, (const uint8_t *)src_d + done, now, cudaMemcpyDeviceToHost, ctx->stream));
      sync(ctx);
      if (fwrite(stage.p, 1, now, f) != now) io_fail("write to", name);
      done += now;
    }
  }

  void take(psc_context *ctx, long id, long beg, long m, const uint32_t *sa_d, const uint32_t *gap_d,
            const uint32_t *gt_fwd_d) override {
    EmBlock &b = blocks[id];
    b.beg = beg;
    b.m = m;
    if (gt_fwd_d && getenv("PSC_KEEP_PARTS")) {
      // the block's own gt_begin bits as the reference stores them: reversed, LSB first, one file per block
      // registered for the reversed index range [n - end, n - beg) (partial_sufsort.hpp:231-234, 356-358)
      const long nbytes = (m + 7) / 8;
      DevBuf<uint8_t> rev(ctx, nbytes);
      bits_forward_to_rev(ctx, gt_fwd_d, m, rev.p, beg);
      b.gt_file = out_path + "." + random_hash();
      FILE *f = fopen(b.gt_file.c_str(), "w");
      if (!f) io_fail("cannot create", b.gt_file);
      write_from_device(ctx, rev.p, (size_t)nbytes, f, b.gt_file);
      if (fclose(f) != 0) io_fail("close", b.gt_file);
    }
    // partial suffix array: block-relative int32 items (distributed_file<int>, psascan.hpp:117-120)
    b.sa_base = out_path + ".distrfile." + random_hash();
    b.items_per_part = std::max(1L, part_bytes / 4);
    b.n_parts = (int)div_up(m, b.items_per_part);
    for (int k = 0; k < b.n_parts; ++k) {
      const std::string name = b.sa_base + ".part" + std::to_string(k);
      FILE *f = fopen(name.c_str(), "w");
      if (!f) io_fail("cannot create", name);
      const long first = (long)k * b.items_per_part, cnt = std::min(b.items_per_part, m - first);
      write_from_device(ctx, sa_d + first, (size_t)cnt * 4, f, name);
      if (fclose(f) != 0) io_fail("close", name);
    }
    if (!gap_d) return;
    // gap array: coarse carries stay in memory, the counters go to the v-byte file
    const long cnt = m + 1, tiles = coarse_carry_tiles(cnt);
    DevBuf<uint64_t> carry_d(ctx, tiles + 1);
    coarse_carries_u32(ctx, gap_d, cnt, carry_d.p);
    b.carry.resize(tiles + 1);
    d2h(ctx, b.carry.data(), carry_d.p, (size_t)tiles + 1);
    sync(ctx);
    const uint64_t tail = (uint64_t)(n - (beg + m));
    if (b.carry[tiles] != tail) {
      set_error("gap counters of block %ld add up to %llu, its tail has %llu suffixes: a 32-bit counter overflowed", id,
                (unsigned long long)b.carry[tiles], (unsigned long long)tail);
      throw CudaError{PSC_ELIMIT};
    }
    DevBuf<uint8_t> bytes;
    const size_t nb = vbyte_encode_d(ctx, gap_d, cnt, bytes, &b.voff);
    b.gap_file = gap_prefix + ".gap." + random_hash();
    FILE *f = fopen(b.gap_file.c_str(), "w");
    if (!f) io_fail("cannot create", b.gap_file);
    write_from_device(ctx, bytes.p, nb, f, b.gap_file);
    if (fclose(f) != 0) io_fail("close", b.gap_file);
  }

  // ---- split of one boundary in block k: the largest u with u + E[u] <= q (merge_split_kernel, on the host) -------
  void split(EmBlock &b, uint64_t q, uint64_t *u_out, uint64_t *e_out) {
    const long cnt = b.m + 1, tiles = (long)b.carry.size() - 1;
    long lo = 0, hi = tiles - 1;
    while (lo < hi) {
      const long mid = (lo + hi + 1) >> 1;
      if ((uint64_t)(mid * COARSE_TILE) + b.carry[mid] <= q) lo = mid; else hi = mid - 1;
    }
    // the counters of tile `lo` from the file
    const size_t nbytes = (size_t)(b.voff[lo + 1] - b.voff[lo]);
    std::vector<uint8_t> raw(nbytes);
    FILE *f = fopen(b.gap_file.c_str(), "r");
    if (!f) io_fail("cannot open", b.gap_file);
    if (fseeko(f, (off_t)b.voff[lo], SEEK_SET) != 0 || fread(raw.data(), 1, nbytes, f) != nbytes) io_fail("read", b.gap_file);
    fclose(f);
    long k = lo * COARSE_TILE;
    uint64_t e = b.carry[lo];
    const long stop = std::min(cnt, k + COARSE_TILE) - 1;
    size_t pos = 0;
    while (k < stop) {
      uint64_t g = 0;
      int shift = 0;
      while (raw[pos] & 0x80) { g |= (uint64_t)(raw[pos++] & 0x7f) << shift; shift += 7; }
      g |= (uint64_t)raw[pos++] << shift;
      if ((uint64_t)(k + 1) + e + g > q) break;
      e += g;
      ++k;
    }
    *u_out = (uint64_t)k;
    *e_out = e;
  }

  // file -> device bytes, through the pinned buffer
  void read_to_device(psc_context *ctx, FILE *f, const std::string &name, void *dst_d, size_t bytes) {
    for (size_t done = 0; done < bytes;) {
      const size_t now = std::min(stage.bytes, bytes - done);
      if (fread(stage.p, 1, now, f) != now) io_fail("read", name);
      PSC_CUDA(cudaMemcpyAsync((uint8_t *)dst_d + done, stage.p, now, cudaMemcpyHostToDevice, ctx->stream));
      sync(ctx);
      done += now;
    }
  }

  // the next cnt SA entries of block b (parts are read front to back and deleted when exhausted, like
  // distributed_file::close_and_destroy_cur_file, io/distributed_file.hpp:218-234)
  void read_sa(psc_context *ctx, EmBlock &b, long cnt, uint32_t *dst_d, bool keep) {
    while (cnt > 0) {
      const int part = (int)(b.sa_next / b.items_per_part);
      const std::string name = b.sa_base + ".part" + std::to_string(part);
      if (part != b.sa_part) {
        if (b.sa_fp) { fclose(b.sa_fp); if (!keep) unlink((b.sa_base + ".part" + std::to_string(b.sa_part)).c_str()); }
        b.sa_fp = fopen(name.c_str(), "r");
        if (!b.sa_fp) io_fail("cannot open", name);
        b.sa_part = part;
      }
      const long in_part = b.sa_next - (long)part * b.items_per_part;
      const long now = std::min(cnt, b.items_per_part - in_part);
      read_to_device(ctx, b.sa_fp, name, dst_d, (size_t)now * 4);
      dst_d += now;
      b.sa_next += now;
      cnt -= now;
    }
  }

  void merge(psc_context *ctx, const char *out_file) {
    PhaseTimer timer(ctx, &ctx->stats.ms_merge);
    const int B = (int)blocks.size();
    const bool keep = getenv("PSC_KEEP_PARTS") != nullptr;
    const long C = merge_chunk_entries();
    const int nb = (int)div_up(n, C) + 1;
    MergePlan plan;
    plan.resize(B, nb);
    for (int i = 0; i < nb; ++i) plan.q[i] = (uint64_t)std::min(n, (long)i * C);
    for (int k = 0; k < B; ++k)
      for (int i = 0; i < nb; ++i) {
        const uint64_t q = plan.q[(size_t)k * nb + i];
        uint64_t u = q, e = 0;
        if (k < B - 1) split(blocks[k], q, &u, &e);
        plan.u[(size_t)k * nb + i] = u;
        plan.e[(size_t)k * nb + i] = e;
        if (k + 1 < B) plan.q[(size_t)(k + 1) * nb + i] = q - u;
      }
    FILE *out = fopen(out_file, "w");
    if (!out) io_fail("cannot create", out_file);
    MergeScratch scr;
    scr.reserve(ctx, std::min(n, C) + 16, need_hi_plane(n));
    DevBuf<uint32_t> sa_buf(ctx, std::min(n, C) + 16), gap_buf(ctx, std::min(n, C) + B + 16);
    DevBuf<uint8_t> out5(ctx, (size_t)(std::min(n, C) + 16) * 5);
    std::vector<LevelSlice> lv(B);
    for (int i = 0; i + 1 < nb; ++i) {
      const long a = (long)plan.Q(0, i), e = (long)plan.Q(0, i + 1);
      long sa_at = 0, gap_at = 0;
      for (int k = 0; k < B; ++k) {
        EmBlock &b = blocks[k];
        const long ua = (long)plan.U(k, i), ub = (long)plan.U(k, i + 1), cb = ub - ua;
        if (ua != b.sa_next) { set_error("external merge: block %d is not consumed in order", k); throw CudaError{PSC_EARG}; }
        uint32_t *sa_k = sa_buf.p + sa_at, *gap_k = gap_buf.p + gap_at;
        if (cb > 0) {
          read_sa(ctx, b, cb, sa_k, keep);
          if (k < B - 1) {
            // counters [ua, ub]: the bytes from the cursor to the end of the tile that holds counter ub
            if (!b.gap_fp) { b.gap_fp = fopen(b.gap_file.c_str(), "r"); if (!b.gap_fp) io_fail("cannot open", b.gap_file); }
            const long t_end = std::min((long)b.voff.size() - 1, ub / COARSE_TILE + 1);
            const size_t nbytes = (size_t)(b.voff[t_end] - b.gap_byte);
            DevBuf<uint8_t> raw(ctx, nbytes);
            if (fseeko(b.gap_fp, (off_t)b.gap_byte, SEEK_SET) != 0) io_fail("seek in", b.gap_file);
            read_to_device(ctx, b.gap_fp, b.gap_file, raw.p, nbytes);
            size_t byte_of_next = 0;
            const long got = vbyte_decode_d(ctx, raw.p, (long)nbytes, cb + 1, gap_k, cb, &byte_of_next);
            if (got < cb + 1) { set_error("external merge: gap file %s ends early", b.gap_file.c_str()); throw CudaError{PSC_EARG}; }
            b.gap_byte += byte_of_next;  // counter ub is the first one of the next chunk's slice
          }
        }
        lv[k] = plan.slice(k, i, sa_k, gap_k, b.beg);
        sa_at += (cb + 3) / 4 * 4;
        gap_at += (cb + 1 + 3) / 4 * 4;
      }
      cascade_chunk(ctx, lv.data(), B, e - a, scr, out5.p);
      write_from_device(ctx, out5.p, (size_t)(e - a) * 5, out, out_file);
    }
    if (fclose(out) != 0) io_fail("close", out_file);
    if (keep) write_manifest();
    for (auto &b : blocks) {
      if (b.sa_fp) fclose(b.sa_fp);
      if (b.gap_fp) fclose(b.gap_fp);
      b.sa_fp = b.gap_fp = nullptr;
    }
    if (!keep) remove_files();
  }

  // PSC_KEEP_PARTS: what a reader of the files needs to know (the format test feeds it to the reference's own
  // merge()): beg end sa_base items_per_part n_parts gap_file gt_begin_file
  void write_manifest() {
    const std::string name = out_path + ".parts.txt";
    FILE *f = fopen(name.c_str(), "w");
    if (!f) io_fail("cannot create", name);
    for (auto &b : blocks)
      fprintf(f, "%ld %ld %s %ld %d %s %s\n", b.beg, b.beg + b.m, b.sa_base.c_str(), b.items_per_part, b.n_parts,
              b.gap_file.empty() ? "-" : b.gap_file.c_str(), b.gt_file.empty() ? "-" : b.gt_file.c_str());
    fclose(f);
  }

  void remove_files() {
    for (auto &b : blocks) {
      for (int k = 0; k < b.n_parts; ++k) unlink((b.sa_base + ".part" + std::to_string(k)).c_str());
      if (!b.gap_file.empty()) unlink(b.gap_file.c_str());
      if (!b.gt_file.empty()) unlink(b.gt_file.c_str());
    }
  }
};

}  // namespace psc
using namespace psc;

extern "C" {

int psc_construct_sa_files_d(psc_context *ctx, const uint8_t *text_d, long n, long max_block, const char *out_path,
                             const char *gap_prefix, long part_bytes) {
  EmSink *sink = nullptr;
  try {
    if (!ctx || n <= 0 || !text_d || !out_path || !gap_prefix) { set_error("psc_construct_sa_files_d: bad argument"); return PSC_EARG; }
    PSC_CUDA(cudaSetDevice(ctx->device));
    if (max_block <= 0) max_block = psc_max_block_size_files(ctx, n);
    if (max_block <= 0) { set_error("text does not fit the HBM budget"); return PSC_ENOMEM; }
    max_block = std::min(max_block, (1L << 31) - 4096);
    if (part_bytes <= 0) part_bytes = 1L << 30;
    sink = new EmSink(out_path, gap_prefix, n, div_up(n, max_block), part_bytes);
    construct_sa_with_sink(ctx, text_d, n, max_block, sink);
    sink->merge(ctx, out_path);
    sync(ctx);
    trace_report(ctx);
    delete sink;
    return PSC_OK;
  } catch (const CudaError &e) {
    if (sink) { if (!getenv("PSC_KEEP_PARTS")) sink->remove_files(); delete sink; }
    return e.code;
  }
}

/* bits [r_lo, r_hi) of a bit sequence stored as a multifile (io/multifile.hpp:45-68: file k holds bits [beg[k], end[k])
 * from its own bit 0, LSB first, bitvector.hpp:61-67 / 77-79), packed from bit 0 of out: what
 * io/multifile_bit_stream_reader.hpp:79-93 hands out bit by bit */
int psc_multifile_read_bits(int n_files, const long *beg, const long *end, const char *const *names, long r_lo, long r_hi,
                            uint8_t *out) {
  if (n_files < 0 || r_lo < 0 || r_hi < r_lo || (r_hi > r_lo && (!beg || !end || !names || !out))) {
    set_error("psc_multifile_read_bits: bad argument");
    return PSC_EARG;
  }
  memset(out, 0, (size_t)((r_hi - r_lo + 7) / 8));
  long covered = 0;
  for (int k = 0; k < n_files; ++k) {
    const long lo = std::max(r_lo, beg[k]), hi = std::min(r_hi, end[k]);
    if (hi <= lo) continue;
    FILE *f = fopen(names[k], "r");
    if (!f) { set_error("cannot open %s: %s", names[k], strerror(errno)); return PSC_EARG; }
    const long b0 = (lo - beg[k]) >> 3, b1 = (hi - beg[k] + 7) >> 3;
    std::vector<uint8_t> raw((size_t)(b1 - b0));
    if (fseeko(f, (off_t)b0, SEEK_SET) != 0 || fread(raw.data(), 1, raw.size(), f) != raw.size()) {
      fclose(f);
      set_error("cannot read %s", names[k]);
      return PSC_EARG;
    }
    fclose(f);
    for (long r = lo; r < hi; ++r) {
      const long q = r - beg[k] - (b0 << 3);
      if ((raw[q >> 3] >> (q & 7)) & 1) out[(r - r_lo) >> 3] |= (uint8_t)(1u << ((r - r_lo) & 7));
    }
    covered += hi - lo;
  }
  if (covered != r_hi - r_lo) { set_error("psc_multifile_read_bits: the files do not cover [%ld, %ld)", r_lo, r_hi); return PSC_EARG; }
  return PSC_OK;
}

int psc_vbyte_encode(psc_context *ctx, const uint32_t *vals, long count, uint8_t *out, long *nbytes) {
  try {
    if (!ctx || count < 0 || !nbytes || (count > 0 && (!vals || !out))) { set_error("psc_vbyte_encode: bad argument"); return PSC_EARG; }
    *nbytes = 0;
    if (count == 0) return PSC_OK;
    PSC_CUDA(cudaSetDevice(ctx->device));
    DevBuf<uint32_t> v(ctx, count);
    h2d(ctx, v.p, vals, count);
    DevBuf<uint8_t> bytes;
    const size_t nb = vbyte_encode_d(ctx, v.p, count, bytes, nullptr);
    d2h(ctx, out, bytes.p, nb);
    sync(ctx);
    *nbytes = (long)nb;
    return PSC_OK;
  } catch (const CudaError &e) { return e.code; }
}

int psc_vbyte_decode(psc_context *ctx, const uint8_t *bytes, long nbytes, uint32_t *vals, long cap, long *count) {
  try {
    if (!ctx || nbytes < 0 || cap < 0 || !count || (nbytes > 0 && (!bytes || !vals))) { set_error("psc_vbyte_decode: bad argument"); return PSC_EARG; }
    *count = 0;
    if (nbytes == 0) return PSC_OK;
    PSC_CUDA(cudaSetDevice(ctx->device));
    DevBuf<uint8_t> b(ctx, nbytes);
    DevBuf<uint32_t> v(ctx, std::max(1L, cap));
    h2d(ctx, b.p, bytes, nbytes);
    const long got = vbyte_decode_d(ctx, b.p, nbytes, cap, v.p, -1, nullptr);
    d2h(ctx, vals, v.p, (size_t)std::min(got, cap));
    sync(ctx);
    *count = got;
    return PSC_OK;
  } catch (const CudaError &e) { return e.code; }
}

}  // extern "C"
