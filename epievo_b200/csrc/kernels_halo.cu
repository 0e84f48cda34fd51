// kernels_halo.cu -- pack / unpack of the per-colour halo records (multi-GPU).
//
// After a colour pass, each GPU sends the ONE boundary site of that colour on each side
// (rows of every branch, state bits, and the three tri_llh entries the update wrote) to
// the neighbour, which stores it in its halo store.  Radius-2 dependence + stride-3
// colouring mean exactly one of the three boundary sites per side can have changed.
#include "kernels.h"
#include "rows.cuh"

namespace epv {

namespace {

struct RecLayout {
  int sw, cw, n_br, K;
  __host__ __device__ int state_at() const { return 4; }
  __host__ __device__ int cnt_at() const { return 4 + sw; }
  __host__ __device__ int jmp_at() const { return 4 + sw + cw; }
  __host__ __device__ size_t words() const { return (size_t)4 + sw + cw + (size_t)n_br*K; }
};

__host__ __device__ inline RecLayout
layout_of(int n_nodes, int halo_k) {
  RecLayout L;
  L.n_br = n_nodes - 1; L.K = halo_k;
  int W = 1;
  while (W*8 < n_nodes) W *= 2;                 // the context's padded state width
  L.sw = (W + 7)/8;
  L.cw = (L.n_br + 3)/4;
  return L;
}

__device__ void
pack_one(const HaloArgs &a, int64_t site, unsigned long long *rec) {
  const RecLayout L = layout_of(a.pv.n_br + 1, a.pv.halo_k);
  const uint64_t s = (uint64_t)site;
  if (threadIdx.x == 0) {
    double *d = reinterpret_cast<double *>(rec);
    d[0] = s >= 1 ? a.tri_llh[s - 1] : 0.0;
    d[1] = a.tri_llh[s];
    d[2] = s + 1 < a.n_sites ? a.tri_llh[s + 1] : 0.0;
    rec[3] = (unsigned long long)(a.global0 + site);
    uint8_t *sb = reinterpret_cast<uint8_t *>(rec + L.state_at());
    for (int w = 0; w < L.sw*8; ++w) sb[w] = w < a.W ? a.state[s*a.W + w] : 0;
  }
  uint16_t *cnt = reinterpret_cast<uint16_t *>(rec + L.cnt_at());
  double *jmp = reinterpret_cast<double *>(rec + L.jmp_at());
  for (int bi = threadIdx.x; bi < L.n_br; bi += blockDim.x) {
    const Row r = load_row(a.pv, a.state, a.W, bi, s);
    int n = r.n;
    atomicMax(a.hmax, (unsigned)n);
    if (n > L.K) { atomicOr(a.err, ERR_HALO_CAPACITY); n = 0; }
    cnt[bi] = (uint16_t)n;
    for (int k = 0; k < n; ++k) jmp[(size_t)bi*L.K + k] = r.ptr[k];
  }
}

// The same record taken BEFORE the CSR rebuild of the pass: a.pv is the CSR the pass read; an accepted
// update's rows are in the staging blob (u16 counts per branch, then the jump times in branch order).
__device__ void
pack_one_staged(const HaloArgs &a, int64_t site, unsigned long long *rec) {
  const uint32_t so = a.stage_off[(uint64_t)site/3];
  if (so == STG_REJECTED) { pack_one(a, site, rec); return; }
  const RecLayout L = layout_of(a.pv.n_br + 1, a.pv.halo_k);
  const uint64_t s = (uint64_t)site;
  if (threadIdx.x == 0) {
    double *d = reinterpret_cast<double *>(rec);
    d[0] = s >= 1 ? a.tri_llh[s - 1] : 0.0;
    d[1] = a.tri_llh[s];
    d[2] = s + 1 < a.n_sites ? a.tri_llh[s + 1] : 0.0;
    rec[3] = (unsigned long long)(a.global0 + site);
    uint8_t *sb = reinterpret_cast<uint8_t *>(rec + L.state_at());
    for (int w = 0; w < L.sw*8; ++w) sb[w] = w < a.W ? a.state[s*a.W + w] : 0;
  }
  uint16_t *cnt = reinterpret_cast<uint16_t *>(rec + L.cnt_at());
  double *jmp = reinterpret_cast<double *>(rec + L.jmp_at());
  if (so == STG_EMPTY) {
    for (int bi = threadIdx.x; bi < L.n_br; bi += blockDim.x) cnt[bi] = 0;
    return;
  }
  const int hw = (L.n_br + 3) >> 2;
  const uint16_t *pn = reinterpret_cast<const uint16_t *>(a.stage + so);
  const double *pj = a.stage + so + hw;
  for (int bi = threadIdx.x; bi < L.n_br; bi += blockDim.x) {
    int n = pn[bi], before = 0;
    for (int k = 0; k < bi; ++k) before += pn[k];
    atomicMax(a.hmax, (unsigned)n);
    if (n > L.K) { atomicOr(a.err, ERR_HALO_CAPACITY); n = 0; }
    cnt[bi] = (uint16_t)n;
    for (int k = 0; k < n; ++k) jmp[(size_t)bi*L.K + k] = pj[before + k];
  }
}

template <bool STAGED>
__global__ void __launch_bounds__(128)
halo_pack_kernel(const HaloArgs a, int64_t site_lo, unsigned long long *rec_lo,
                 int64_t site_hi, unsigned long long *rec_hi) {
  if (blockIdx.x == 0) { if (site_lo >= 0) { if (STAGED) pack_one_staged(a, site_lo, rec_lo); else pack_one(a, site_lo, rec_lo); } }
  else if (site_hi >= 0) { if (STAGED) pack_one_staged(a, site_hi, rec_hi); else pack_one(a, site_hi, rec_hi); }
}

__device__ void
unpack_one(const HaloArgs &a, int64_t site, const unsigned long long *rec) {
  const RecLayout L = layout_of(a.pv.n_br + 1, a.pv.halo_k);
  const uint64_t s = (uint64_t)site;
  if (rec[3] != (unsigned long long)(a.global0 + site)) {
    if (threadIdx.x == 0) atomicOr(a.err, ERR_HALO_MISMATCH);
    return;
  }
  if (threadIdx.x == 0) {
    const double *d = reinterpret_cast<const double *>(rec);
    if (s >= 1) a.tri_llh[s - 1] = d[0];
    a.tri_llh[s] = d[1];
    if (s + 1 < a.n_sites) a.tri_llh[s + 1] = d[2];
    const uint8_t *sb = reinterpret_cast<const uint8_t *>(rec + L.state_at());
    for (int w = 0; w < a.W; ++w) a.state[s*a.W + w] = sb[w];
  }
  const int slot = s < a.pv.own_lo ? (int)s : HALO_W + (int)(s - a.pv.own_hi);
  const uint16_t *cnt = reinterpret_cast<const uint16_t *>(rec + L.cnt_at());
  const double *jmp = reinterpret_cast<const double *>(rec + L.jmp_at());
  for (int bi = threadIdx.x; bi < L.n_br; bi += blockDim.x) {
    const int n = cnt[bi];
    a.hcnt[slot*L.n_br + bi] = (uint16_t)n;
    for (int k = 0; k < n; ++k) a.hjmp[(size_t)(slot*L.n_br + bi)*L.K + k] = jmp[(size_t)bi*L.K + k];
  }
}

__global__ void __launch_bounds__(128)
halo_unpack_kernel(const HaloArgs a, int64_t site_lo, const unsigned long long *rec_lo,
                   int64_t site_hi, const unsigned long long *rec_hi) {
  if (blockIdx.x == 0) { if (site_lo >= 0) unpack_one(a, site_lo, rec_lo); }
  else if (site_hi >= 0) unpack_one(a, site_hi, rec_hi);
}

}  // namespace

size_t halo_record_words(int n_nodes, int halo_k) { return layout_of(n_nodes, halo_k).words(); }

void
launch_halo_pack(const HaloArgs &a, int64_t site_lo, unsigned long long *rec_lo, int64_t site_hi,
                 unsigned long long *rec_hi, cudaStream_t st, bool staged) {
  if (staged) halo_pack_kernel<true><<<2, 128, 0, st>>>(a, site_lo, rec_lo, site_hi, rec_hi);
  else halo_pack_kernel<false><<<2, 128, 0, st>>>(a, site_lo, rec_lo, site_hi, rec_hi);
}

void
launch_halo_unpack(const HaloArgs &a, int64_t site_lo, const unsigned long long *rec_lo,
                   int64_t site_hi, const unsigned long long *rec_hi, cudaStream_t st) {
  halo_unpack_kernel<<<2, 128, 0, st>>>(a, site_lo, rec_lo, site_hi, rec_hi);
}

}  // namespace epv
