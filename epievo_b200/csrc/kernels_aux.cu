// kernels_aux.cu -- tri_llh cache (reset), deterministic per-site intermediates for the
// parity tests, and a driver for the end-conditioned sampler.
#include "kernels.h"
#include "mh_site2.cuh"

namespace epv {

namespace {

// SingleSiteSampler::reset (SingleSiteSampler.cpp:470-474): tri_llh of interior sites.
__global__ void __launch_bounds__(256)
tri_llh_kernel(const ModelParams mp, const PathView pv, const uint8_t *state, int W,
               const TreeDev tree, uint64_t n_sites, double *tri_llh) {
  __shared__ ModelParams s_mp;
  if (threadIdx.x == 0) s_mp = mp;
  __syncthreads();
  const uint64_t s = (uint64_t)blockIdx.x*blockDim.x + threadIdx.x;
  if (s >= n_sites) return;
  double x = 0.0;
  if (s >= 1 && s + 1 < n_sites) {
    SiteCtx c; c.pv = pv; c.state = state; c.W = W; c.tree = tree; c.site = s;
    x = stored_tri_llh(s_mp, c);
  }
  tri_llh[s] = x;
}

// Single thread; uses the sampler's own merge / transition / pruning routines and
// additionally materialises what the sampler streams (a2, a3, a4, a7).
__global__ void
site_debug_kernel(const ModelParams mp, const PathView pv, const uint8_t *state, int W,
                  const TreeDev tree, uint64_t site, const DebugOut out) {
  __shared__ ModelParams s_mp;
  s_mp = mp;
  SiteCtx c; c.pv = pv; c.state = state; c.W = W; c.tree = tree; c.site = site;
  double *q = out.q;                       // [2*n_nodes], global scratch doubles as output
  upward(s_mp, c, q);
  unsigned long long k = 0;
  out.seg_off[0] = 0; out.seg_off[1] = 0;
  double logp = 0.0;
  bool overflow = false;
  for (int v = 1; v < tree.n_nodes && !overflow; ++v) {
    const Row L = load_row(pv, state, W, v - 1, site - 1);
    const Row R = load_row(pv, state, W, v - 1, site + 1);
    const Row M = load_row(pv, state, W, v - 1, site);
    const double tot = tree.branch[v];
    const int n_int = L.n + R.n + 1;
    if (k + n_int > out.cap) { overflow = true; break; }
    // p[i] for every interval by the backward recursion (process_branch_above)
    {
      RevMerge m; m.start(L, R, tot);
      double p0 = q[2*v], p1 = q[2*v + 1];
      for (int i = n_int - 1; i >= 0; --i) {
        double len; int ctx; bool first;
        m.next(len, ctx, first);
        const P2 P = trans_prob(s_mp.rate[ctx], s_mp.rate[ctx | 2], len);
        const double a = P.p00*p0 + P.p01*p1, b = P.p10*p0 + P.p11*p1;
        p0 = a; p1 = b;
        out.p[2*(k + i)] = p0; out.p[2*(k + i) + 1] = p1;
      }
    }
    // forward walk: interval list (collect_segment_info) and the proposal
    // log-probability of the current path (proposal_prob_branch, :272-312)
    FwdMerge m; m.start(L, R, tot);
    int start = M.init, end = M.init, jend = 0, jstart = 0;
    double end_time = 0.0;
    for (int i = 0; i < n_int; ++i) {
      double len; int ctx; bool last;
      m.next(len, ctx, last);
      out.rate0[k + i] = s_mp.rate[ctx]; out.rate1[k + i] = s_mp.rate[ctx | 2];
      out.trip0[k + i] = ctx; out.trip1[k + i] = ctx | 2; out.len[k + i] = len;
      end_time += len;
      while (jend < M.n && M.ptr[jend] < end_time) ++jend;
      if ((jend - jstart) & 1) end ^= 1;
      const P2 P = trans_prob(s_mp.rate[ctx], s_mp.rate[ctx | 2], len);
      const double PT0 = start ? P.p10 : P.p00;
      const double Pse = start ? (end ? P.p11 : P.p10) : (end ? P.p01 : P.p00);
      logp -= log(Pse);
      const double nx0 = (i == n_int - 1) ? q[2*v] : out.p[2*(k + i + 1)];
      const double p0 = PT0/out.p[2*(k + i) + start]*nx0;
      logp += end == 0 ? log(p0) : log(1.0 - p0);
      jstart = jend; start = end;
    }
    k += n_int;
    out.seg_off[v + 1] = k;
  }
  const int ls = state_bit(state, W, site - 1, 1), rs = state_bit(state, W, site + 1, 1);
  const double a0 = (s_mp.T[ls*2 + 0]*s_mp.T[0*2 + rs])*q[0];
  const double a1 = (s_mp.T[ls*2 + 1]*s_mp.T[1*2 + rs])*q[1];
  *out.root_p0 = a0/(a0 + a1);
  *out.logp = logp;
  *out.n_intervals = overflow ? ~0ull : k;
}

__global__ void __launch_bounds__(128)
end_cond_kernel(double r0, double r1, int a, int b, double T, uint64_t seed, uint64_t n,
                uint64_t cap, uint32_t *n_jumps, double *jumps) {
  const uint64_t i = (uint64_t)blockIdx.x*blockDim.x + threadIdx.x;
  if (i >= n) return;
  Philox rng; rng.init(seed, 0u, i, 1u);
  double buf[64];
  const int c = cap < 64 ? (int)cap : 64;
  int k = end_cond_sample(r0, r1, a, b, T, 0.0, rng, buf, 0, 64);
  n_jumps[i] = k < 0 ? 0xffffffffu : (uint32_t)k;
  for (int j = 0; j < c; ++j) jumps[i*cap + j] = j < k ? buf[j] : 0.0;
}

}  // namespace

void
launch_tri_llh(const ModelParams &mp, const PathView &pv, const uint8_t *state, int W,
               const TreeDev &tree, uint64_t n_sites, double *tri_llh, cudaStream_t st) {
  const unsigned blocks = (unsigned)((n_sites + 255)/256);
  tri_llh_kernel<<<blocks, 256, 0, st>>>(mp, pv, state, W, tree, n_sites, tri_llh);
}

void
launch_site_debug(const ModelParams &mp, const PathView &pv, const uint8_t *state, int W,
                  const TreeDev &tree, uint64_t site, const DebugOut &out,
                  cudaStream_t st) {
  site_debug_kernel<<<1, 1, 0, st>>>(mp, pv, state, W, tree, site, out);
}

void
launch_end_cond(double r0, double r1, int a, int b, double T, uint64_t seed, uint64_t n,
                uint64_t cap, uint32_t *n_jumps, double *jumps, cudaStream_t st) {
  end_cond_kernel<<<(unsigned)((n + 127)/128), 128, 0, st>>>(r0, r1, a, b, T, seed, n, cap,
                                                             n_jumps, jumps);
}

}  // namespace epv
