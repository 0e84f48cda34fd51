// mh_site.cuh -- one Metropolis-Hastings update of one site's whole-tree history given
// both neighbours (SingleSiteSampler::Metropolis_Hastings_site,
// SingleSiteSampler.cpp:482-536), as a per-thread device routine.
//
// Per-thread storage is O(n_nodes): Felsenstein q per node and the proposal's jump
// count per node; interval lists are never materialised -- the neighbour merge is
// re-walked (forward for sampling, backward for pruning) straight from the CSR rows.
#pragma once
#include "rows.cuh"
#include "endcond.cuh"

namespace epv {

constexpr int PBUF = 8;   // per-branch cache of p vectors for the downward pass

struct SiteCtx {
  PathView pv;
  const uint8_t *state;
  int W;
  TreeDev tree;
  uint64_t site;          // local site index
};

// Felsenstein upward pruning (pruning, SingleSiteSampler.cpp:80-157).  q[2*node+k]:
// leaves get the indicator of the site's current end state (:122-126); an internal
// node's q is the product of its children's p at the top of their branches
// (:128-131), accumulated as the children finish (reverse preorder).  Along a branch
// p[i] = P_i p[i+1], the last interval using q (:91-106).
__device__ __forceinline__ void
upward(const ModelParams &mp, const SiteCtx &c, double *q) {
  const int n_nodes = c.tree.n_nodes;
  for (int v = 0; v < n_nodes; ++v) {
    double q0 = 1.0, q1 = 1.0;
    if (c.tree.subtree[v] == 1) {
      const uint32_t *o = c.pv.off + (uint64_t)(v - 1)*c.pv.off_pitch + (c.site - c.pv.own_lo);
      const int leaf = state_bit(c.state, c.W, c.site, v) ^ ((o[1] - o[0]) & 1);
      q0 = leaf ? 0.0 : 1.0; q1 = leaf ? 1.0 : 0.0;
    }
    q[2*v] = q0; q[2*v + 1] = q1;
  }
  for (int v = n_nodes - 1; v > 0; --v) {
    const Row L = load_row(c.pv, c.state, c.W, v - 1, c.site - 1);
    const Row R = load_row(c.pv, c.state, c.W, v - 1, c.site + 1);
    RevMerge m; m.start(L, R, c.tree.branch[v]);
    double p0 = q[2*v], p1 = q[2*v + 1];
    bool first = false;
    while (!first) {
      double len; int ctx;
      m.next(len, ctx, first);
      const P2 P = trans_prob(mp.rate[ctx], mp.rate[ctx | 2], len);
      const double a = P.p00*p0 + P.p01*p1, b = P.p10*p0 + P.p11*p1;
      p0 = a; p1 = b;
    }
    const int par = c.tree.parent[v];
    q[2*par] *= p0; q[2*par + 1] *= p1;
  }
}

// Fills pbuf[2*i..] with p[i+1] (the "next" vector of interval i) for i < PBUF of
// branch v by a backward walk from q[v]; returns the interval count.
__device__ __forceinline__ int
branch_next_vectors(const ModelParams &mp, const Row &L, const Row &R, double tot,
                    double q0, double q1, double *pbuf) {
  const int n_int = L.n + R.n + 1;
  RevMerge m; m.start(L, R, tot);
  double p0 = q0, p1 = q1;
  for (int i = n_int - 1; i >= 0; --i) {
    if (i < PBUF) { pbuf[2*i] = p0; pbuf[2*i + 1] = p1; }
    if (i == 0) break;
    double len; int ctx; bool first;
    m.next(len, ctx, first);
    const P2 P = trans_prob(mp.rate[ctx], mp.rate[ctx | 2], len);
    const double a = P.p00*p0 + P.p01*p1, b = P.p10*p0 + P.p11*p1;
    p0 = a; p1 = b;
  }
  return n_int;
}

// "next" vector of interval i >= PBUF, recomputed from q (rare: > PBUF intervals).
static __device__ __noinline__ void
next_vector_slow(const ModelParams &mp, const Row &L, const Row &R, double tot,
                 double q0, double q1, int i, double &n0, double &n1) {
  const int n_int = L.n + R.n + 1;
  RevMerge m; m.start(L, R, tot);
  double p0 = q0, p1 = q1;
  for (int k = n_int - 1; k > i; --k) {
    double len; int ctx; bool first;
    m.next(len, ctx, first);
    const P2 P = trans_prob(mp.rate[ctx], mp.rate[ctx | 2], len);
    const double a = P.p00*p0 + P.p01*p1, b = P.p10*p0 + P.p11*p1;
    p0 = a; p1 = b;
  }
  n0 = p0; n1 = p1;
}

}  // namespace epv
