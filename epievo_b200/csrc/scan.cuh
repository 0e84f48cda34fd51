// scan.cuh -- single-pass chained scan (decoupled look-back) across tiles, shared by the
// CSR rebuild and the compact import.  One u64 descriptor per (branch, tile): value in the
// low 32 bits, FLAG_AGG (tile aggregate) / FLAG_INC (inclusive prefix) above, and an epoch
// tag in the top bits so the array is never cleared between launches.  Tiles take their
// index from an atomic ticket, so tile order is scheduling order and a tile only ever
// waits for tiles that already run.
#pragma once
#include "common.cuh"

namespace epv {

constexpr unsigned long long FLAG_AGG = 1ull << 32, FLAG_INC = 2ull << 32;

__device__ __forceinline__ unsigned long long
ld_desc(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void
st_desc(unsigned long long *p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}

// Exclusive prefix of this tile's total for one branch; executed by a full warp.
__device__ __forceinline__ uint32_t
tile_base(unsigned long long *desc, uint32_t tile, uint32_t epoch, uint32_t total,
          int lane) {
  const unsigned long long tag = (unsigned long long)epoch << 34;
  if (tile == 0) {
    if (lane == 0) st_desc(desc, tag | FLAG_INC | total);
    return 0;
  }
  if (lane == 0) st_desc(desc + tile, tag | FLAG_AGG | total);
  uint32_t base = 0;
  int64_t look = (int64_t)tile - 1;
  for (;;) {
    const int64_t idx = look - lane;
    unsigned long long d;
    bool ready;
    for (unsigned ns = 32;; ns = ns < 1024 ? 2*ns : ns) {
      d = idx >= 0 ? ld_desc(desc + idx) : (tag | FLAG_INC);
      ready = (d >> 34) == epoch && (d & (FLAG_AGG | FLAG_INC)) != 0;
      if (!__any_sync(0xffffffffu, !ready)) break;
      __nanosleep(ns);            // back off: a spinning warp must not steal issue slots
    }
    const unsigned inc = __ballot_sync(0xffffffffu, (d & FLAG_INC) != 0);
    const int first = inc ? __ffs(inc) - 1 : 32;
    uint32_t v = lane <= first ? (uint32_t)d : 0u;
    v = warp_sum(v);
    base += v;
    if (inc) break;
    look -= 32;
  }
  if (lane == 0) st_desc(desc + tile, tag | FLAG_INC | (base + total));
  return base;
}

}  // namespace epv
