// kernels_indep.cu -- the site-independent E-step of epievo_initialization (SURVEY 8f-3;
// IndepSite.cpp:53-260, :266-303).  Every site evolves as a plain 2-state CTMC with rates
// (r0, r1) on every branch, so one table of per-branch 2x2 matrices serves all sites:
//   indep_expect_kernel  expectation_sufficient_statistics (:221-239): Felsenstein up,
//                        joint end-state posteriors down, analytic E[J], E[D] per branch
//   indep_sample_kernel  update_paths_indep (:243-260): the same upward pass, then end
//                        states and end-conditioned jump times sampled root to leaves; run
//                        as a colour pass so that the staging blob + csr_rebuild_kernel of
//                        the MH sweep install the new rows
//   indep_stats_kernel   compute_sufficient_statistics (:266-303): J / D per state from the
//                        stored paths
#include "endcond.cuh"
#include "kernels.h"
#include "mh_site2.cuh"

namespace epv {

namespace {

constexpr uint32_t STG_REJECTED = 0xffffffffu, STG_EMPTY = 0xfffffffeu;
constexpr int IN_THREADS = 128;

// Leaf indicator / product of the children's p (upward_process, IndepSite.cpp:53-89).
// q[2v..2v+1]; a leaf's end state is its row's init state flipped once per jump.
template <int MAXN>
__device__ __forceinline__ void
indep_upward(const PathView &pv, const uint8_t *state, int W, const TreeDev &tree, const IndepTab *tab,
             uint64_t site, double *q) {
  const int n = tree.n_nodes;
  for (int v = 0; v < n; ++v) { q[2*v] = 1.0; q[2*v + 1] = 1.0; }
  for (int v = n - 1; v >= 1; --v) {
    if (tree.subtree[v] == 1) {
      const Row r = load_row(pv, state, W, v - 1, site);
      const int leaf = r.init ^ (r.n & 1);
      q[2*v] = leaf ? 0.0 : 1.0; q[2*v + 1] = leaf ? 1.0 : 0.0;
    }
    const double *P = tab[v].P;
    const double p0 = P[0]*q[2*v] + P[1]*q[2*v + 1], p1 = P[2]*q[2*v] + P[3]*q[2*v + 1];
    const int par = tree.parent[v];
    q[2*par] *= p0; q[2*par + 1] *= p1;
  }
}

template <int MAXN>
__global__ void __launch_bounds__(IN_THREADS)
indep_expect_kernel(const PathView pv, const uint8_t *state, int W, const TreeDev tree,
                    const IndepTab *tab, double pi0, double *acc) {
  extern __shared__ double s_acc[];               // [n_nodes*4]: J0 J1 D0 D1 per node
  const int n = tree.n_nodes, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 4*n; i += IN_THREADS) s_acc[i] = 0.0;
  __syncthreads();
  double q[2*MAXN], p0m[MAXN];
  const uint64_t n_own = pv.own_hi - pv.own_lo;
  const uint64_t per = (uint64_t)gridDim.x*IN_THREADS;
  for (uint64_t base = (uint64_t)blockIdx.x*IN_THREADS; base < n_own; base += per) {
    const uint64_t i = base + threadIdx.x;
    const bool on = i < n_own;
    if (on) {
      indep_upward<MAXN>(pv, state, W, tree, tab, pv.own_lo + i, q);
      const double a = pi0*q[0], b = (1.0 - pi0)*q[1];       // root_post_prob0, :92-99
      p0m[0] = a/(a + b);
    }
    for (int v = 1; v < n; ++v) {
      double j0 = 0.0, j1 = 0.0, d0 = 0.0, d1 = 0.0;
      if (on) {
        const IndepTab &t = tab[v];
        const double q0 = q[2*v], q1 = q[2*v + 1];
        const double p_0 = t.P[0]*q0 + t.P[1]*q1, p_1 = t.P[2]*q0 + t.P[3]*q1;
        const double pu = p0m[tree.parent[v]];
        double pj[4] = {t.P[0]*q0*pu/p_0, t.P[1]*q1*pu/p_0,              // joint_post, :103-120
                        t.P[2]*q0*(1.0 - pu)/p_1, t.P[3]*q1*(1.0 - pu)/p_1};
        const double Z = pj[0] + pj[1] + pj[2] + pj[3];
#pragma unroll
        for (int k = 0; k < 4; ++k) pj[k] /= Z;
        p0m[v] = pj[0] + pj[2];
#pragma unroll
        for (int k = 0; k < 4; ++k) {                                     // :146-153
          j0 += pj[k]*t.J0[k]; j1 += pj[k]*t.J1[k]; d0 += pj[k]*t.D0[k]; d1 += pj[k]*t.D1[k];
        }
      }
      j0 = warp_sum(j0); j1 = warp_sum(j1); d0 = warp_sum(d0); d1 = warp_sum(d1);
      if (lane == 0) {
        atomicAdd(&s_acc[4*v], j0); atomicAdd(&s_acc[4*v + 1], j1);
        atomicAdd(&s_acc[4*v + 2], d0); atomicAdd(&s_acc[4*v + 3], d1);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 4*n; i += IN_THREADS) if (s_acc[i] != 0.0) atomicAdd(acc + i, s_acc[i]);
}

// J[state] / D[state] of every branch from the stored rows.
__global__ void __launch_bounds__(IN_THREADS)
indep_stats_kernel(const PathView pv, const uint8_t *state, int W, const TreeDev tree, double *acc) {
  extern __shared__ double s_acc[];
  const int n = tree.n_nodes, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 4*n; i += IN_THREADS) s_acc[i] = 0.0;
  __syncthreads();
  const uint64_t n_own = pv.own_hi - pv.own_lo;
  const uint64_t per = (uint64_t)gridDim.x*IN_THREADS;
  for (uint64_t base = (uint64_t)blockIdx.x*IN_THREADS; base < n_own; base += per) {
    const uint64_t i = base + threadIdx.x;
    const bool on = i < n_own;
    for (int v = 1; v < n; ++v) {
      double J[2] = {0.0, 0.0}, D[2] = {0.0, 0.0};
      if (on) {
        const Row r = load_row(pv, state, W, v - 1, pv.own_lo + i);
        int prev = r.init;
        double time = 0.0;
        for (int k = 0; k < r.n; ++k) {
          const double t = r.ptr[k];
          J[prev] += 1.0; D[prev] += t - time;
          prev ^= 1; time = t;
        }
        D[prev] += tree.branch[v] - time;
      }
      const double j0 = warp_sum(J[0]), j1 = warp_sum(J[1]), d0 = warp_sum(D[0]), d1 = warp_sum(D[1]);
      if (lane == 0) {
        atomicAdd(&s_acc[4*v], j0); atomicAdd(&s_acc[4*v + 1], j1);
        atomicAdd(&s_acc[4*v + 2], d0); atomicAdd(&s_acc[4*v + 3], d1);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 4*n; i += IN_THREADS) if (s_acc[i] != 0.0) atomicAdd(acc + i, s_acc[i]);
}

// sampling_downward (IndepSite.cpp:176-214) for the sites of one colour; the new rows go
// to the staging blob exactly as an accepted MH proposal does.  Philox streams: node 1 =
// the end-state draws, node 2 + v = the jump times of branch v (so a rerun with a larger
// buffer reproduces the same path).
template <int MAXN, int MAXJ>
__global__ void __launch_bounds__(IN_THREADS)
indep_sample_kernel(const PassArgs a, const IndepTab *tab, double r0, double r1) {
  const int t = threadIdx.x, n = a.tree.n_nodes, n_br = n - 1;
  const uint64_t g = (uint64_t)blockIdx.x*IN_THREADS + t;
  if (g >= a.n_groups) return;
  const uint64_t st = 3*g + a.colour;
  const int64_t gsite = a.global0 + (int64_t)st;
  uint32_t so = STG_REJECTED;
  if (st >= a.in.own_lo && st < a.in.own_hi) {
    double q[2*MAXN], pj_local[MAXJ];
    uint16_t pn[MAXN];
    uint8_t pend[MAXN];
    indep_upward<MAXN>(a.in, a.state, a.W, a.tree, tab, st, q);
    double *pj = pj_local;
    int cap = MAXJ, total = 0;
    pend[0] = (uint8_t)state_bit(a.state, a.W, st, 1);       // the root keeps its state (:186)
    for (int attempt = 0; attempt < 2; ++attempt) {
      Philox rng; rng.init(a.seed, a.sweep, (uint64_t)gsite, 1u);
      total = 0;
      for (int v = 1; v < n && total >= 0; ++v) {
        const int start = pend[a.tree.parent[v]];
        const double *P = tab[v].P;
        const double q0 = q[2*v], q1 = q[2*v + 1];
        const double p_start = P[2*start]*q0 + P[2*start + 1]*q1;
        const double p0 = P[2*start]*q0/p_start;                        // :198-200
        const int e = rng.uniform() > p0;
        Philox rj; rj.init(a.seed, a.sweep, (uint64_t)gsite, 2u + (uint32_t)v);
        const int before = total;
        total = end_cond_sample(r0, r1, start, e, a.tree.branch[v], 0.0, rj, pj, total, cap);
        if (total >= 0) {
          if (total - before > 65535) total = -1;
          else { pn[v] = (uint16_t)(total - before); pend[v] = (uint8_t)e; }
        }
      }
      if (total != -1 || attempt == 1) break;
      const uint32_t slot = atomicAdd(a.scratch_ctr, 1u);
      if (slot >= a.scratch_slots) { atomicOr(a.err, ERR_SCRATCH_SLOTS); break; }
      pj = a.scratch + (uint64_t)slot*a.scratch_cap; cap = (int)a.scratch_cap;
    }
    if (total == -1) atomicOr(a.err, ERR_SCRATCH_SIZE);
    if (total == -2) atomicOr(a.err, ERR_SAMPLER_STUCK);
    if (total >= 0) {
      // state bits: row v starts in its parent's end state; bit 0 (the empty root row) takes the
      // root state as sampled_path.front() does (:186)
      for (int w = 0; w < a.W; ++w) {
        unsigned sb = 0;
        for (int k = 0; k < 8; ++k) {
          const int v = w*8 + k;
          if (v < n) sb |= (unsigned)(v == 0 ? pend[0] : pend[a.tree.parent[v]]) << k;
        }
        a.state[st*a.W + w] = (uint8_t)sb;
      }
      if (total == 0) so = STG_EMPTY;
      else {
        const int hw = (n_br + 3) >> 2;
        const unsigned long long need = (unsigned long long)hw + total;
        const unsigned long long at = atomicAdd(a.stage_ctr, need);
        if (at + need > a.stage_cap) atomicOr(a.err, ERR_STAGE_CAPACITY);
        else {
          so = (uint32_t)at;
          double *blob = a.stage + at;
          uint16_t *cnt = reinterpret_cast<uint16_t *>(blob);
          for (int bi = 0; bi < 4*hw; ++bi) cnt[bi] = bi < n_br ? pn[bi + 1] : 0;
          for (int k = 0; k < total; ++k) blob[hw + k] = pj[k];
        }
      }
    }
  }
  a.stage_off[g] = so;
}

template <int MAXN>
void expect_t(const PathView &pv, const uint8_t *state, int W, const TreeDev &tree, const IndepTab *tab,
              double pi0, double *acc, unsigned blocks, size_t smem, cudaStream_t st) {
  indep_expect_kernel<MAXN><<<blocks, IN_THREADS, smem, st>>>(pv, state, W, tree, tab, pi0, acc);
}

}  // namespace

void
launch_indep_expect(const PathView &pv, const uint8_t *state, int W, const TreeDev &tree,
                    const IndepTab *tab, double pi0, double *acc, int sm_count, cudaStream_t st) {
  const int n = tree.n_nodes;
  const uint64_t n_own = pv.own_hi - pv.own_lo;
  const size_t smem = (size_t)4*n*sizeof(double);
  unsigned blocks = (unsigned)std::min<uint64_t>((n_own + IN_THREADS - 1)/IN_THREADS, (uint64_t)sm_count*8);
  if (!blocks) blocks = 1;
  cudaMemsetAsync(acc, 0, smem, st);
  if (n <= 8) expect_t<8>(pv, state, W, tree, tab, pi0, acc, blocks, smem, st);
  else if (n <= 16) expect_t<16>(pv, state, W, tree, tab, pi0, acc, blocks, smem, st);
  else if (n <= 64) expect_t<64>(pv, state, W, tree, tab, pi0, acc, blocks, smem, st);
  else expect_t<256>(pv, state, W, tree, tab, pi0, acc, blocks, smem, st);
}

void
launch_indep_stats(const PathView &pv, const uint8_t *state, int W, const TreeDev &tree, double *acc,
                   int sm_count, cudaStream_t st) {
  const uint64_t n_own = pv.own_hi - pv.own_lo;
  const size_t smem = (size_t)4*tree.n_nodes*sizeof(double);
  unsigned blocks = (unsigned)std::min<uint64_t>((n_own + IN_THREADS - 1)/IN_THREADS, (uint64_t)sm_count*8);
  if (!blocks) blocks = 1;
  cudaMemsetAsync(acc, 0, smem, st);
  indep_stats_kernel<<<blocks, IN_THREADS, smem, st>>>(pv, state, W, tree, acc);
}

void
launch_indep_sample(const PassArgs &a, const IndepTab *tab, double r0, double r1, cudaStream_t st) {
  const unsigned blocks = (unsigned)((a.n_groups + IN_THREADS - 1)/IN_THREADS);
  const int n = a.tree.n_nodes;
  if (n <= 8) indep_sample_kernel<8, 16><<<blocks, IN_THREADS, 0, st>>>(a, tab, r0, r1);
  else if (n <= 16) indep_sample_kernel<16, 32><<<blocks, IN_THREADS, 0, st>>>(a, tab, r0, r1);
  else if (n <= 64) indep_sample_kernel<64, 64><<<blocks, IN_THREADS, 0, st>>>(a, tab, r0, r1);
  else indep_sample_kernel<256, 128><<<blocks, IN_THREADS, 0, st>>>(a, tab, r0, r1);
}

}  // namespace epv
