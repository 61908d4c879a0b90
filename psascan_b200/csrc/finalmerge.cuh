// finalmerge.cuh -- the N-way merge by output range (finalmerge.cu), shared by api.cu and dist.cu.
#pragma once
#include <vector>
#include "common.cuh"

namespace psc {

// one sorted block kept until the final merge
struct MergePiece {
  long beg = 0, m = 0;
  const uint32_t *sa = nullptr;   // m block-relative entries
  const uint32_t *gap = nullptr;  // m + 1 counters; null for the rightmost block
};

// what one chunk of output slots needs from one block: cb suffixes and the gap counters in front of them
struct LevelSlice {
  const uint32_t *sa = nullptr;   // block-relative SA entries [ua, ua + cb)
  const uint32_t *gap = nullptr;  // gap[ua ..] (cb entries are used; one more may be read)
  long cb = 0;
  uint64_t skip = 0;  // entries of the first run that lie before the chunk
  long beg = 0;       // text position of the block
};

struct MergeScratch {
  DevBuf<uint64_t> E;  // scanned gap slice of one level
  long cap = 0;
  void reserve(psc_context *ctx, long entries, bool hi_plane /* unused: values are 40 bits wide on the way out */);
};

// host copy of the split tables: for level k and boundary i, Q = slot in M_k, U = block-k suffixes below
// it, Ev = E_k[U]
struct MergePlan {
  int B = 0, nb = 0;
  std::vector<uint64_t> q, u, e;
  void resize(int levels, int boundaries);
  uint64_t Q(int k, int i) const { return q[(size_t)k * nb + i]; }
  uint64_t U(int k, int i) const { return u[(size_t)k * nb + i]; }
  uint64_t Ev(int k, int i) const { return e[(size_t)k * nb + i]; }
  LevelSlice slice(int k, int i, const uint32_t *sa_at_ua, const uint32_t *gap_at_ua, long beg) const;
};

// where finished chunks go when the output is not one resident device buffer (host staging)
struct ChunkSink {
  virtual uint8_t *buffer(psc_context *ctx, int chunk, long entries) const = 0;  // device buffer to pack into
  virtual void done(psc_context *ctx, int chunk, long first_entry, long entries) const = 0;
  virtual ~ChunkSink() {}
};

// takes a finished block out of the pipeline (external-memory mode, em.cu): its SA and gap array are released by
// the pipeline right after the call
struct PieceSink {
  // gt_fwd_d: the gt bits relative to the block begin, bit j - 1 <-> text position j (null for the leftmost block)
  virtual void take(psc_context *ctx, long id, long beg, long m, const uint32_t *sa_d, const uint32_t *gap_d /* null: rightmost block */,
                    const uint32_t *gt_fwd_d) = 0;
  virtual ~PieceSink() {}
};
// the block pipeline of api.cu with every block handed to `pieces` instead of being kept for the resident merge
void construct_sa_with_sink(psc_context *ctx, const uint8_t *text_d, long n, long max_block, PieceSink *pieces);

long merge_chunk_entries();
void merge_split_slice(psc_context *ctx, const uint32_t *gap_slice_d, long cnt, long u_off, uint64_t e_base,
                       const uint64_t *carry_d, bool last_slice, const uint64_t *q_d, int nq, uint64_t *u_d,
                       uint64_t *e_d);
void merge_split_next(psc_context *ctx, const uint64_t *q_d, const uint64_t *u_d, int nq, uint64_t *q_next_d);
void cascade_chunk(psc_context *ctx, const LevelSlice *lv, int n_levels, long total, MergeScratch &scr,
                   uint8_t *out5_d);
void final_merge_resident(psc_context *ctx, const std::vector<MergePiece> &pc, long n, uint8_t *out5_d,
                          const ChunkSink *sink = nullptr);

}  // namespace psc
