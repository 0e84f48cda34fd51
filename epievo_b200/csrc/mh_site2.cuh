// mh_site2.cuh -- downward sampling, proposal likelihoods and the acceptance test.
#pragma once
#include "mh_site.cuh"

namespace epv {

// A proposed whole-tree history of one site: jump times of all branches packed in
// preorder in pj[0..total), pn[v] jumps on branch v, pend[v] = end state of branch v
// (pend[0] = root state).  Branch v starts in pend[parent[v]].
struct Proposal {
  double *pj;
  int cap;
  uint16_t *pn;
  uint8_t *pend;
};

// Downward sampling (downward_sampling, SingleSiteSampler.cpp:227-255; per branch
// :180-221): the root keeps its state (SAMPLE_ROOT is always false, :441); per
// interval the end state is drawn with p0 = P(prev->0) next[0] / p[i][prev] (:199-206)
// and the jumps inside from the end-conditioned law (:209-213).
// Returns the total jump count, or <0 (-1: pj too small, -2: sampler stuck).
__device__ __forceinline__ int
downward(const ModelParams &mp, const SiteCtx &c, const double *q, uint64_t seed,
         uint32_t sweep, uint64_t gsite, Proposal &pr) {
  const int n_nodes = c.tree.n_nodes;
  pr.pend[0] = (uint8_t)state_bit(c.state, c.W, c.site, 1);
  pr.pn[0] = 0;
  int total = 0;
  for (int v = 1; v < n_nodes; ++v) {
    const Row L = load_row(c.pv, c.state, c.W, v - 1, c.site - 1);
    const Row R = load_row(c.pv, c.state, c.W, v - 1, c.site + 1);
    const double tot = c.tree.branch[v];
    double pbuf[2*PBUF];
    const int n_int = branch_next_vectors(mp, L, R, tot, q[2*v], q[2*v + 1], pbuf);
    Philox rng; rng.init(seed, sweep, gsite, (uint32_t)v);
    FwdMerge m; m.start(L, R, tot);
    int prev = pr.pend[c.tree.parent[v]];
    double passed = 0.0;
    const int n_before = total;
    for (int i = 0; i < n_int; ++i) {
      double len; int ctx; bool last;
      m.next(len, ctx, last);
      const double r0 = mp.rate[ctx], r1 = mp.rate[ctx | 2];
      const P2 P = trans_prob(r0, r1, len);
      double nx0, nx1;
      if (i < PBUF) { nx0 = pbuf[2*i]; nx1 = pbuf[2*i + 1]; }
      else next_vector_slow(mp, L, R, tot, q[2*v], q[2*v + 1], i, nx0, nx1);
      const double a0 = (prev ? P.p10 : P.p00)*nx0;
      const double a1 = (prev ? P.p11 : P.p01)*nx1;
      const double p0 = a0/(a0 + a1);
      const int s = rng.uniform() > p0;
      total = end_cond_sample(r0, r1, prev, s, len, passed, rng, pr.pj, total, pr.cap);
      if (total < 0) return total;
      passed += len;
      prev = s;
    }
    if (total - n_before > 65535) return -1;
    pr.pn[v] = (uint16_t)(total - n_before);
    pr.pend[v] = (uint8_t)prev;
  }
  return total;
}

// The three triple log-likelihoods that involve the proposal (log_accept_rate,
// SingleSiteSampler.cpp:422-429; path_log_likelihood :374-391).  The root term is the
// raw probability T(l,m)T(m,r), not its log, as in the reference (:383-384).
__device__ __forceinline__ void
proposal_llh(const ModelParams &mp, const SiteCtx &c, const Proposal &pr,
             bool do_l, bool do_r, double &ll, double &lm, double &lr) {
  const int n_nodes = c.tree.n_nodes;
  const uint64_t s = c.site;
  const int sm = pr.pend[0];
  const int sl1 = state_bit(c.state, c.W, s - 1, 1), sr1 = state_bit(c.state, c.W, s + 1, 1);
  lm = mp.T[sl1*2 + sm]*mp.T[sm*2 + sr1];
  if (do_l) ll = mp.T[state_bit(c.state, c.W, s - 2, 1)*2 + sl1]*mp.T[sl1*2 + sm];
  if (do_r) lr = mp.T[sm*2 + sr1]*mp.T[sr1*2 + state_bit(c.state, c.W, s + 2, 1)];
  int off = 0;
  for (int v = 1; v < n_nodes; ++v) {
    const double tot = c.tree.branch[v];
    Row P; P.ptr = pr.pj + off; P.n = pr.pn[v]; P.init = pr.pend[c.tree.parent[v]];
    off += P.n;
    const Row B = load_row(c.pv, c.state, c.W, v - 1, s - 1);
    const Row C = load_row(c.pv, c.state, c.W, v - 1, s + 1);
    lm += branch_llh(mp, B, P, C, tot);
    if (do_l) ll += branch_llh(mp, load_row(c.pv, c.state, c.W, v - 1, s - 2), B, P, tot);
    if (do_r) lr += branch_llh(mp, P, C, load_row(c.pv, c.state, c.W, v - 1, s + 2), tot);
  }
}

// tri_llh of a stored site (SingleSiteSampler::reset, SingleSiteSampler.cpp:470-474).
__device__ __forceinline__ double
stored_tri_llh(const ModelParams &mp, const SiteCtx &c) {
  const uint64_t s = c.site;
  const int a = state_bit(c.state, c.W, s - 1, 1), b = state_bit(c.state, c.W, s, 1),
    d = state_bit(c.state, c.W, s + 1, 1);
  double x = mp.T[a*2 + b]*mp.T[b*2 + d];
  for (int v = 1; v < c.tree.n_nodes; ++v)
    x += branch_llh(mp, load_row(c.pv, c.state, c.W, v - 1, s - 1),
                    load_row(c.pv, c.state, c.W, v - 1, s),
                    load_row(c.pv, c.state, c.W, v - 1, s + 1), c.tree.branch[v]);
  return x;
}

// Whole update.  Returns 1 accepted, 0 rejected, <0 capacity/sampler failure (nothing
// written).  On accept, state bits and tri_llh are updated in place (race-free within
// a stride-3 colour class); the caller installs the new rows.
// The proposal log-probabilities of old and new path cancel exactly in the
// Hastings ratio (both equal -log q_root[root state] for any leaf-consistent path;
// SURVEY 8a-7), so  llr = (llh_l + llh_r) - (llh_l_orig + llh_r_orig)  (:417-432).
__device__ __forceinline__ int
mh_site(const ModelParams &mp, const SiteCtx &c, uint8_t *state_rw, double *tri_llh,
        uint64_t seed, uint32_t sweep, uint64_t gsite, uint64_t n_global,
        double *q, Proposal &pr, int &total_out) {
  upward(mp, c, q);
  const int total = downward(mp, c, q, seed, sweep, gsite, pr);
  if (total < 0) return total;
  total_out = total;
  const bool do_l = gsite > 1, do_r = gsite + 2 < n_global;
  const double l_old = tri_llh[c.site - 1], r_old = tri_llh[c.site + 1];
  double ll = l_old, lm = 0.0, lr = r_old;
  proposal_llh(mp, c, pr, do_l, do_r, ll, lm, lr);
  const double llr = (ll + lr) - l_old - r_old;
  Philox rng; rng.init(seed, sweep, gsite, 0u);
  const double u = rng.uniform();
  if (!(llr >= 0.0 || u < exp(llr))) return 0;
  tri_llh[c.site - 1] = ll; tri_llh[c.site] = lm; tri_llh[c.site + 1] = lr;
  for (int w = 0; w < c.W; ++w) {
    unsigned bits = 0;
    for (int k = 0; k < 8; ++k) {
      const int v = w*8 + k;
      if (v >= 1 && v < c.tree.n_nodes) bits |= (unsigned)pr.pend[c.tree.parent[v]] << k;
    }
    state_rw[c.site*c.W + w] = (uint8_t)bits;
  }
  return 1;
}

}  // namespace epv
