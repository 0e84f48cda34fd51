// common.cuh -- shared device/host definitions for the epievo_b200 kernels (sm_100a).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace epv {

constexpr int N_TRIP = 8;

// ---- error flags raised by kernels (sticky, read back by the host) ---------------
enum : unsigned {
  ERR_JUMP_CAPACITY = 1u,   // a branch's jump array would overflow
  ERR_SCRATCH_SLOTS = 2u,   // more long proposals in one pass than scratch slots
  ERR_SCRATCH_SIZE  = 4u,   // one proposal longer than a scratch slot
  ERR_SAMPLER_STUCK = 8u,   // end-conditioned sampler exceeded its iteration cap
  ERR_BAD_INPUT     = 16u,
  ERR_HALO_CAPACITY = 32u,  // a boundary row has more jumps than a halo row holds
  ERR_HALO_MISMATCH = 64u,  // a halo record arrived for another site / colour
  ERR_STAGE_CAPACITY = 128u,// accepted proposals of one pass exceed the staging buffer
};

// ---- model constants, one copy in __constant__ memory per context upload ---------
struct ModelParams {
  double rate[N_TRIP];      // triplet rates, index l*4+m*2+r (epievo_utils.hpp:85-88)
  double log_rate[N_TRIP];  // SingleSiteSampler::log_rates (SingleSiteSampler.cpp:465-468)
  double T[4];              // horizontal transition probs, row-major
};

// ---- device view of the path storage ------------------------------------------------
// Per branch (node b >= 1, index bi = b-1) a CSR over the local sites:
//   off[bi*off_pitch + s]   uint32, s in [0, n_sites]   offsets into that branch's jumps
//   jmp + jbase[bi]         double                       jump times, strictly inside (0, T_b)
// Initial states are site-major bit sets: bit (node&7) of state[s*W + (node>>3)].
// Only the OWNED sites [own_lo, own_hi) of the local index space live in the CSR (row
// index = site - own_lo).  The <= 3 halo sites on either side (copies of the neighbour
// GPUs' boundary sites, refreshed after every colour pass) are kept in a small
// fixed-capacity halo store so that a refresh never forces a re-compaction:
//   hcnt[slot*n_br + bi], hjmp[(slot*n_br + bi)*halo_k ...], slot = site for the low
//   halo, 3 + (site - own_hi) for the high halo.
constexpr int HALO_W = 3;        // halo width in sites (dependence radius is 2)
constexpr int HALO_K_MIN = 48;   // smallest jump capacity of one halo row; the context's halo_k is sized from
                                 // the boundary rows at upload and regrown between sweeps (overflow = loud error)

struct PathView {
  const uint32_t *off;
  const double *jmp;
  const uint64_t *jbase;    // [n_branches] start of each branch's slice of jmp
  uint64_t off_pitch;
  uint64_t own_lo, own_hi;  // owned local sites [own_lo, own_hi) of n_local
  uint64_t n_local;
  const uint16_t *hcnt;
  const double *hjmp;
  int n_br;
  int halo_k;               // jump capacity of one halo row
};

struct TreeDev {
  int n_nodes;
  const int *parent;        // [n_nodes]
  const int *subtree;       // [n_nodes]
  const double *branch;     // [n_nodes] current branch lengths == rows' tot_time
};

// ---- Philox4x32-10 (Salmon et al., SC'11), counter-based ---------------------------
struct Philox {
  uint32_t c0, c1, c2, c3;  // counter
  uint32_t k0, k1;          // key
  double spare;
  int have_spare;

  // key = seed; counter = (site_lo, site_hi:8 | node<<8, sweep, draw index)
  __device__ __forceinline__ void
  init(uint64_t seed, uint32_t sweep, uint64_t site, uint32_t node) {
    k0 = (uint32_t)seed; k1 = (uint32_t)(seed >> 32);
    c0 = (uint32_t)site;
    c1 = ((uint32_t)(site >> 32) & 0xffu) | (node << 8);
    c2 = sweep;
    c3 = 0;
    have_spare = 0;
  }

  __device__ __forceinline__ void
  block(uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3) const {
    uint32_t x0 = c0, x1 = c1, x2 = c2, x3 = c3, a = k0, b = k1;
#pragma unroll
    for (int i = 0; i < 10; ++i) {
      const uint32_t hi0 = __umulhi(0xD2511F53u, x0), lo0 = 0xD2511F53u*x0;
      const uint32_t hi1 = __umulhi(0xCD9E8D57u, x2), lo1 = 0xCD9E8D57u*x2;
      const uint32_t y0 = hi1 ^ x1 ^ a, y2 = hi0 ^ x3 ^ b;
      x0 = y0; x1 = lo1; x2 = y2; x3 = lo0;
      a += 0x9E3779B9u; b += 0xBB67AE85u;
    }
    r0 = x0; r1 = x1; r2 = x2; r3 = x3;
  }

  // uniform on the open interval (0,1), 52 random bits: (k + 1/2) * 2^-52
  static __device__ __forceinline__ double
  to_unit(uint32_t hi, uint32_t lo) {
    const uint64_t k = (((uint64_t)hi << 32) | lo) >> 12;
    return ((double)k + 0.5)*2.220446049250313e-16;
  }

  // two uniforms from one fresh block (keeps the lanes of a warp converged: no
  // data-dependent "spare" state)
  __device__ __forceinline__ void
  uniform2(double &u1, double &u2) {
    uint32_t r0, r1, r2, r3;
    block(r0, r1, r2, r3);
    ++c3;
    u1 = to_unit(r0, r1); u2 = to_unit(r2, r3);
  }

  __device__ __forceinline__ double
  uniform() {
    if (have_spare) { have_spare = 0; return spare; }
    uint32_t r0, r1, r2, r3;
    block(r0, r1, r2, r3);
    ++c3;
    spare = to_unit(r2, r3);
    have_spare = 1;
    return to_unit(r0, r1);
  }
};

// ---- small helpers -----------------------------------------------------------------------
__device__ __forceinline__ int
state_bit(const uint8_t *state, int W, uint64_t site, int node) {
  return (state[site*W + (node >> 3)] >> (node & 7)) & 1;
}

// All state bits of one site as 64-bit words (W is a power of two; W <= 8: one load).
__device__ __forceinline__ uint64_t
state_word(const uint8_t *state, int W, uint64_t site, int word = 0) {
  const uint8_t *p = state + site*W;
  switch (W) {
  case 1: return *p;
  case 2: return *reinterpret_cast<const uint16_t *>(p);
  case 4: return *reinterpret_cast<const uint32_t *>(p);
  default: return reinterpret_cast<const uint64_t *>(p)[word];
  }
}

template <typename T>
__device__ __forceinline__ T
warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace epv
