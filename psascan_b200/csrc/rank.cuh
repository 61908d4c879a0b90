// rank.cuh -- device-side query of the one-sector rank layout (see rank.cu).
#pragma once
#include "common.cuh"

namespace psc {

constexpr int RK_BITS = 224;  // symbols per sector: 7 words of bits + 1 word of count = 32 bytes

struct RankView {
  const uint32_t *sectors;  // [n_codes][n_sectors][8]
  long n_sectors;
  uint32_t m;
};

// occ(code, i) = number of positions p < i whose BWT symbol has this dense code.
// One 32-byte sector is read per query (a single 256-bit load, sm_100a LDG.E.256).  `.L2::64B`:
// without it the L2 fetches 128 bytes from HBM per miss (measured, profiles/r1_summary.md section 4);
// the rate of such loads is the same, but the unused half of the traffic is what a sort running on
// the other stream can use.
__device__ __forceinline__ uint32_t rank_occ(const RankView &rv, int code, uint32_t i) {
  uint32_t s = i / RK_BITS;
  uint32_t r = i - s * RK_BITS;
  const uint32_t *p = rv.sectors + ((size_t)code * (size_t)rv.n_sectors + s) * 8;
  uint32_t w[8];
  asm volatile("ld.global.nc.L1::no_allocate.L2::64B.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]),
                 "=r"(w[7])
               : "l"(p));
  uint32_t cnt = w[0];
#pragma unroll
  for (int k = 0; k < 7; ++k) {
    int nb = (int)r - 32 * k;
    nb = nb < 0 ? 0 : (nb > 32 ? 32 : nb);
    uint32_t mask = nb == 32 ? 0xffffffffu : ((1u << nb) - 1u);
    cnt += __popc(w[k + 1] & mask);
  }
  return cnt;
}

}  // namespace psc
