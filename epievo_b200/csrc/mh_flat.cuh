// mh_flat.cuh -- the Metropolis-Hastings site update restructured for SIMT.
//
// Same mathematics as mh_site (mh_site2.cuh; SingleSiteSampler.cpp:482-536), different
// control flow.  The nested "for branch { for interval { ... } }" loops, whose trip
// counts differ between the 32 sites of a warp, are flattened into single loops whose
// iteration is one unit of work (one neighbour interval, one jump event), so a warp
// runs max-over-lanes of the per-site TOTAL instead of the sum over branches of the
// per-branch maximum:
//   1. upward pruning: one interval per iteration, backward in time, branches in reverse
//      preorder; partial vectors p[i] (and exp terms h_i) of multi-interval branches go
//      to a small per-thread stack that the downward pass pops in exactly reverse order;
//   2. downward pass: one interval per iteration; draws the interval's end state and,
//      when start == end, the "no jump at all" Bernoulli; intervals that need jump times
//      are queued (FIFO), not sampled inline;
//   3. the queue is drained with end_cond_sample (rare work, done with the lanes that
//      have it);
//   4. the three triple log-likelihoods of the proposal come from ONE five-way merge per
//      branch over the rows of sites s-2..s+2, one event per iteration.
// Transition probabilities use per-context tables (pi0 = r0/s, pi1 = r1/s, s): no fp64
// division; the end-state and no-jump draws compare products instead of ratios.
#pragma once
#include "mh_site2.cuh"

namespace epv {

struct CtxTables {          // indexed by ctx = l*4 + r (mid bit clear): 0, 1, 4, 5
  double s[8], pi0[8], pi1[8];
};

__device__ __forceinline__ void
fill_tables(const ModelParams &mp, CtxTables &tb, int t) {
  if (t < 8) {
    const int ctx = t & ~2;
    const double r0 = mp.rate[ctx], r1 = mp.rate[ctx | 2], s = r0 + r1;
    tb.s[t] = s; tb.pi0[t] = r0/s; tb.pi1[t] = r1/s;
  }
}

// Init-state bit sets of sites s-2..s+2, read once per update (WORDS x 64 node bits).
template <int WORDS>
struct StateWin {
  uint64_t w[5][WORDS];
  __device__ __forceinline__ void
  load(const uint8_t *state, int W, uint64_t s, uint64_t n_sites) {
#pragma unroll
    for (int k = 0; k < 5; ++k) {
#pragma unroll
      for (int j = 0; j < WORDS; ++j) w[k][j] = 0;
      const int64_t site = (int64_t)s + k - 2;
      if (site >= 0 && site < (int64_t)n_sites)
#pragma unroll
        for (int j = 0; j < WORDS; ++j)
          if (j*8 < W) w[k][j] = state_word(state, W, (uint64_t)site, j);
    }
  }
  // d = site offset relative to s (-2..2)
  __device__ __forceinline__ int bit(int d, int node) const {
    return (int)((w[d + 2][WORDS == 1 ? 0 : node >> 6] >> (node & 63)) & 1u);
  }
};

// Jump times and count of one row; HALO=false skips the halo-store branch (single-GPU
// contexts own every local site).
template <bool HALO>
__device__ __forceinline__ void
row_data(const PathView &pv, int bi, uint64_t site, const double *&ptr, int &n) {
  if (!HALO || (site >= pv.own_lo && site < pv.own_hi)) {
    const uint32_t *o = pv.off + (uint64_t)bi*pv.off_pitch + (site - pv.own_lo);
    const uint32_t a = o[0], b = o[1];
    ptr = pv.jmp + pv.jbase[bi] + a; n = (int)(b - a);
  }
  else {
    const int slot = site < pv.own_lo ? (int)site : HALO_W + (int)(site - pv.own_hi);
    n = pv.hcnt[slot*pv.n_br + bi];
    ptr = pv.hjmp + (size_t)(slot*pv.n_br + bi)*pv.halo_k;
  }
}

// Rows of sites s-1, s, s+1 on branch bi (own row: count only).
template <bool HALO, typename SW>
__device__ __forceinline__ void
load_lmr(const PathView &pv, const SW &sw, int bi, uint64_t site,
         Row &L, Row &R, int &own_n) {
  L.init = sw.bit(-1, bi + 1);
  R.init = sw.bit(1, bi + 1);
  if (!HALO || (site - 1 >= pv.own_lo && site + 1 < pv.own_hi)) {
    const uint32_t *o = pv.off + (uint64_t)bi*pv.off_pitch + (site - 1 - pv.own_lo);
    const uint32_t o0 = o[0], o1 = o[1], o2 = o[2], o3 = o[3];
    const double *base = pv.jmp + pv.jbase[bi];
    L.ptr = base + o0; L.n = (int)(o1 - o0);
    own_n = (int)(o2 - o1);
    R.ptr = base + o2; R.n = (int)(o3 - o2);
  }
  else {
    const double *dummy;
    row_data<true>(pv, bi, site - 1, L.ptr, L.n);
    row_data<true>(pv, bi, site + 1, R.ptr, R.n);
    row_data<true>(pv, bi, site, dummy, own_n);
  }
}

// One queued interval that needs jump times; lives in SHARED memory so that any lane of
// the warp can sample it.  Pending: a = start time within the branch, b = length,
// meta = v | ctx<<8 | cur<<11 | end<<12 | (per-site item number)<<16.  Done: a, b = the
// (at most two) jump times, meta = v | count<<16 | Q_DONE.
struct QSlot { double a, b; unsigned meta; unsigned pad; };
constexpr unsigned Q_DONE = 1u << 31;

template <int MAXN, int PB, int QCAP>
struct FlatScratch {
  double q[2*MAXN];
  double hb[MAXN];
  double stk[3*PB];
  QSlot *queue;        // this thread's QCAP slots in shared memory
  int *excl;           // this warp's 32 prefix counts in shared memory
};

// Philox stream of the k-th queued interval of a site: node index 2 + k (0 = accept draw,
// 1 = downward pass), so the result does not depend on which lane samples it.
__device__ __forceinline__ int
sample_slot(const ModelParams &mp, const QSlot &it, uint64_t seed, uint32_t sweep, uint64_t gsite,
            double *out, int n, int cap) {
  const int ctx = (it.meta >> 8) & 7, cur = (it.meta >> 11) & 1, end = (it.meta >> 12) & 1;
  Philox rng; rng.init(seed, sweep, gsite, 2u + ((it.meta >> 16) & 0x7fffu));
  return end_cond_sample(mp.rate[ctx], mp.rate[ctx | 2], cur, end, it.b, it.a, rng, out, n, cap,
                         cur == end);
}

// Owner-side collection of the queue (FIFO keeps pj in branch / time order): slots already
// sampled by some lane are copied, the others are sampled here with their own stream.
__device__ __forceinline__ int
collect_queue(const ModelParams &mp, QSlot *queue, int &qn, uint64_t seed, uint32_t sweep,
              uint64_t gsite, Proposal &pr, int total) {
  for (int k = 0; k < qn; ++k) {
    const QSlot it = queue[k];
    const int v = it.meta & 0xff;
    const int before = total;
    if (it.meta & Q_DONE) {
      const int cnt = (it.meta >> 16) & 3;
      if (total + cnt > pr.cap) return -1;
      pr.pj[total] = it.a;
      if (cnt > 1) pr.pj[total + 1] = it.b;
      total += cnt;
    }
    else {
      total = sample_slot(mp, it, seed, sweep, gsite, pr.pj, total, pr.cap);
      if (total < 0) return total;
    }
    const int add = total - before + pr.pn[v];
    if (add > 65535) return -1;
    pr.pn[v] = (uint16_t)add;
  }
  qn = 0;
  return total;
}

// Warp-cooperative sampling of every pending slot of the warp's 32 queues: the slots are
// numbered through a prefix sum of the queue lengths and dealt out 32 at a time, one per
// lane, so the rare jump-time work runs with the lanes full instead of 2-4 active.
// Results with more than two jumps stay pending (the owner samples them in collect_queue
// with the same stream => same result).  Warp-synchronous.
template <int QCAP>
__device__ __forceinline__ void
coop_sample(const ModelParams &mp, QSlot *queue, int *excl, int qn, uint64_t seed, uint32_t sweep,
            uint64_t gsite) {
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  int incl = qn;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int up = __shfl_up_sync(FULL, incl, o);
    if (lane >= o) incl += up;
  }
  const int T = __shfl_sync(FULL, incl, 31);
  excl[lane] = incl - qn;
  __syncwarp();
  QSlot *warp_q = queue - lane*QCAP;                 // lane 0's slots; lane L's are + L*QCAP
  const uint32_t g_lo = (uint32_t)gsite, g_hi = (uint32_t)(gsite >> 32);
  for (int base = 0; base < T; base += 32) {
    const int j = base + lane;
    int owner = 0;
#pragma unroll
    for (int st = 16; st > 0; st >>= 1)
      if (owner + st < 32 && excl[owner + st] <= j) owner += st;
    const uint64_t og = ((uint64_t)__shfl_sync(FULL, g_hi, owner) << 32) | __shfl_sync(FULL, g_lo, owner);
    if (j < T) {
      QSlot *slot = warp_q + owner*QCAP + (j - excl[owner]);
      const QSlot it = *slot;
      double out[2];
      const int n = sample_slot(mp, it, seed, sweep, og, out, 0, 2);
      if (n > 0) {
        slot->a = out[0]; slot->b = n > 1 ? out[1] : 0.0;
        slot->meta = (it.meta & 0xffu) | ((unsigned)n << 16) | Q_DONE;
      }
    }
  }
  __syncwarp();
}

// Returns 1 accepted, 0 rejected (or inactive lane), -1 proposal buffer too small (rerun
// with a larger one: same streams, same proposal), -2 sampler stuck, -3 interval stack too
// small (depends on the neighbours' rows only; the caller falls back to mh_site).
//
// WARP-SYNCHRONOUS: must be called by all 32 lanes of a warp together (lanes without a
// site to update pass active = false).  Every flat loop runs `while (__any_sync(...))`,
// so its trip count is warp-uniform and the lanes reconverge every iteration; there are
// no early returns (a failing lane records its status and idles).  Without this the
// compiler leaves lanes that left a loop at different iterations diverged for the rest
// of the function (measured: 2-4 active lanes per instruction).
template <int MAXN, int PB, int QCAP, bool HALO>
__device__ __forceinline__ int
mh_site_flat(bool active, bool coop, const ModelParams &mp, const CtxTables &tb, const SiteCtx &c,
             uint8_t *state_rw, double *tri_llh, uint64_t seed, uint32_t sweep, uint64_t gsite,
             uint64_t n_global, FlatScratch<MAXN, PB, QCAP> &S, Proposal &pr, int &total_out) {
  constexpr unsigned FULL = 0xffffffffu;
  const int n_nodes = c.tree.n_nodes;
  const uint64_t s = c.site;
  double *q = S.q;
  int status = 1;                       // becomes <0 on failure; the lane then idles
  StateWin<(MAXN + 63)/64> sw;
  if (active) sw.load(c.state, c.W, s, c.pv.n_local);
  // the three passes below walk the CSR offsets of sites s-2..s+3 on every branch: start the
  // loads of all those lines now (no registers held; the lanes of a warp share 3-4 lines per
  // branch) instead of missing on them one branch at a time on the warp's critical path
  if (active && s >= c.pv.own_lo + 2 && s + 3 < c.pv.own_hi) {
    const uint32_t *o = c.pv.off + (s - 2 - c.pv.own_lo);
    for (int bi = n_nodes - 2; bi >= 0; --bi) {
      const uint32_t *p = o + (uint64_t)bi*c.pv.off_pitch;
      asm volatile("prefetch.global.L1 [%0];" :: "l"(p));
    }
    asm volatile("prefetch.global.L2 [%0];" :: "l"(tri_llh + s - 1));
  }

  // ---- 1. upward pruning, flat over (branch, interval) ----------------------------
  // (an INACTIVE lane leaves its outputs alone: the caller re-enters this routine with only the lanes that
  // need a larger proposal buffer active, and the other lanes still hold the result of their first call)
  for (int v = 0; v < n_nodes; ++v) { q[2*v] = 1.0; q[2*v + 1] = 1.0; if (active) pr.pn[v] = 0; }
  int sp = 0;
  {
    int v = active ? n_nodes - 1 : 0;
    RevMerge m;
    double p0 = 0.0, p1 = 0.0;
    bool need_init = true;
    while (__any_sync(FULL, v >= 1)) {
      if (v < 1) continue;
      if (need_init) {
        Row L, R; int own_n;
        load_lmr<HALO>(c.pv, sw, v - 1, s, L, R, own_n);
        m.start(L, R, c.tree.branch[v]);
        if (c.tree.subtree[v] == 1) {
          const int leaf = sw.bit(0, v) ^ (own_n & 1);
          p0 = leaf ? 0.0 : 1.0; p1 = leaf ? 1.0 : 0.0;
          q[2*v] = p0; q[2*v + 1] = p1;
        }
        else { p0 = q[2*v]; p1 = q[2*v + 1]; }
        need_init = false;
      }
      double len; int ctx; bool first;
      m.next(len, ctx, first);
      const double h = exp(-len*tb.s[ctx]);
      const double P00 = tb.pi1[ctx] + tb.pi0[ctx]*h, P11 = tb.pi0[ctx] + tb.pi1[ctx]*h;
      const double a = P00*p0 + (1.0 - P00)*p1, b = (1.0 - P11)*p0 + P11*p1;
      p0 = a; p1 = b;
      if (first) {
        S.hb[v] = h;
        const int par = c.tree.parent[v];
        q[2*par] *= p0; q[2*par + 1] *= p1;
        --v; need_init = true;
      }
      else if (sp >= PB) { status = -3; v = 0; }
      else { S.stk[3*sp] = p0; S.stk[3*sp + 1] = p1; S.stk[3*sp + 2] = h; ++sp; }
    }
  }

  // ---- 2. downward pass, flat over (branch, interval) -----------------------------
  Philox rng; rng.init(seed, sweep, gsite, 1u);
  if (active) pr.pend[0] = (uint8_t)sw.bit(0, 1);
  int total = 0, qn = 0, q_total = 0;
  {
    int v = (active && status > 0) ? 1 : n_nodes, prev = 0, i = 0;
    FwdMerge m;
    double passed = 0.0, h_saved = 0.0;
    bool need_init = true;
    while (__any_sync(FULL, v < n_nodes)) {
      if (v >= n_nodes) continue;
      if (need_init) {
        Row L, R; int own_n;
        load_lmr<HALO>(c.pv, sw, v - 1, s, L, R, own_n);
        m.start(L, R, c.tree.branch[v]);
        prev = pr.pend[c.tree.parent[v]];
        passed = 0.0; i = 0; need_init = false;
      }
      double len; int ctx; bool last;
      m.next(len, ctx, last);
      const double h = i == 0 ? S.hb[v] : h_saved;
      double nx0, nx1;
      if (last) { nx0 = q[2*v]; nx1 = q[2*v + 1]; }
      else { --sp; nx0 = S.stk[3*sp]; nx1 = S.stk[3*sp + 1]; h_saved = S.stk[3*sp + 2]; }
      const double P00 = tb.pi1[ctx] + tb.pi0[ctx]*h, P11 = tb.pi0[ctx] + tb.pi1[ctx]*h;
      const double to0 = prev ? 1.0 - P11 : P00, to1 = prev ? P11 : 1.0 - P00;
      const double a0 = to0*nx0, a1 = to1*nx1;
      double u_end, u_jump;
      rng.uniform2(u_end, u_jump);
      const int e = u_end*(a0 + a1) > a0;                  // end state of the interval
      bool need_jumps = e != prev;
      if (!need_jumps) {
        // P(no jump | start == end) = exp(-r_prev len) / P(prev -> prev)
        // compare u P(prev -> prev) with exp(-x), x = r_prev len, through the alternating-series
        // bounds 1 - x + x^2/2 - x^3/6 <= exp(-x) <= ... + x^4/24; exp itself only in the gap
        // (the decision is the one exp would give: the margin covers the rounding of the bounds)
        const double x = len*(prev ? mp.rate[ctx | 2] : mp.rate[ctx]);
        const double lhs = u_jump*(prev ? P11 : P00);
        const double lo = 1.0 + x*(-1.0 + x*(0.5 - x*(1.0/6.0)));
        const double x2 = x*x;
        if (lhs < lo - 1e-15) need_jumps = false;
        else if (lhs > lo + x2*x2*(1.0/24.0) + 1e-15) need_jumps = true;
        else need_jumps = !(lhs < exp(-x));
      }
      if (need_jumps) {
        if (qn == QCAP) total = collect_queue(mp, S.queue, qn, seed, sweep, gsite, pr, total);
        if (total < 0) { status = total; v = n_nodes; continue; }
        QSlot it; it.a = passed; it.b = len; it.pad = 0;
        it.meta = (unsigned)v | (unsigned)ctx << 8 | (unsigned)prev << 11 | (unsigned)e << 12 |
                  ((unsigned)(q_total++) & 0x7fffu) << 16;
        S.queue[qn++] = it;
      }
      passed += len; prev = e; ++i;
      if (last) { pr.pend[v] = (uint8_t)prev; ++v; need_init = true; }
    }
  }
  // ---- 3. jump times of the queued intervals ---------------------------------------
  if (status <= 0) qn = 0;
  // (warp-uniform switch; without it every owner samples its own queue: same streams, same result)
  if (coop) coop_sample<QCAP>(mp, S.queue, S.excl, qn, seed, sweep, gsite);
  if (status > 0) {
    total = collect_queue(mp, S.queue, qn, seed, sweep, gsite, pr, total);
    if (total < 0) status = total;
  }
  __syncwarp(FULL);
  if (active) total_out = total;

  // ---- 4. the three triple log-likelihoods: one five-way merge per branch ----------
  const bool live = active && status > 0;
  const bool do_l = live && gsite > 1, do_r = live && gsite + 2 < n_global;
  const double l_old = live ? tri_llh[s - 1] : 0.0, r_old = live ? tri_llh[s + 1] : 0.0;
  double xl = 0.0, xm = 0.0, xr = 0.0;
  if (live) {
    const int sm = pr.pend[0];
    const int sl1 = sw.bit(-1, 1), sr1 = sw.bit(1, 1);
    xm = mp.T[sl1*2 + sm]*mp.T[sm*2 + sr1];
    xl = do_l ? mp.T[sw.bit(-2, 1)*2 + sl1]*mp.T[sl1*2 + sm] : l_old;
    xr = do_r ? mp.T[sm*2 + sr1]*mp.T[sr1*2 + sw.bit(2, 1)] : r_old;
  }
  {
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    int v = live ? 1 : n_nodes, poff = 0;
    Row A, B, P, C, E;
    A.n = B.n = P.n = C.n = E.n = 0; A.ptr = B.ptr = P.ptr = C.ptr = E.ptr = pr.pj;
    int iA = 0, iB = 0, iP = 0, iC = 0, iE = 0, bits = 0;   // bits: A<<4 | B<<3 | P<<2 | C<<1 | E
    double prev_l = 0.0, prev_m = 0.0, prev_r = 0.0, tot = 0.0;
    bool need_init = true;
    while (__any_sync(FULL, v < n_nodes)) {
      if (v >= n_nodes) continue;
      if (need_init) {
        const int bi = v - 1;
        row_data<HALO>(c.pv, bi, s - 1, B.ptr, B.n); B.init = sw.bit(-1, v);
        row_data<HALO>(c.pv, bi, s + 1, C.ptr, C.n); C.init = sw.bit(1, v);
        if (do_l) { row_data<HALO>(c.pv, bi, s - 2, A.ptr, A.n); A.init = sw.bit(-2, v); }
        else { A.n = 0; A.init = 0; A.ptr = B.ptr; }
        if (do_r) { row_data<HALO>(c.pv, bi, s + 2, E.ptr, E.n); E.init = sw.bit(2, v); }
        else { E.n = 0; E.init = 0; E.ptr = C.ptr; }
        P.ptr = pr.pj + poff; P.n = pr.pn[v]; P.init = pr.pend[c.tree.parent[v]];
        poff += P.n;
        iA = iB = iP = iC = iE = 0;
        bits = A.init << 4 | B.init << 3 | P.init << 2 | C.init << 1 | E.init;
        prev_l = prev_m = prev_r = 0.0;
        tot = c.tree.branch[v];
        need_init = false;
      }
      // next event: smallest time, ties to the higher site index (the reference's
      // right-before-mid-before-left rule, Path.cpp:275-281, holds for all three triples)
      double t = iE < E.n ? E.ptr[iE] : INF; int k = 0;
      { const double x = iC < C.n ? C.ptr[iC] : INF; if (x < t) { t = x; k = 1; } }
      { const double x = iP < P.n ? P.ptr[iP] : INF; if (x < t) { t = x; k = 2; } }
      { const double x = iB < B.n ? B.ptr[iB] : INF; if (x < t) { t = x; k = 3; } }
      { const double x = iA < A.n ? A.ptr[iA] : INF; if (x < t) { t = x; k = 4; } }
      const bool done = !(t < INF);
      if (done) t = tot;
      const int tl = bits >> 2, tm = (bits >> 1) & 7, tr = bits & 7;
      if (do_l && (k >= 2 || done)) { xl -= mp.rate[tl]*(t - prev_l); prev_l = t; }
      if ((k >= 1 && k <= 3) || done) { xm -= mp.rate[tm]*(t - prev_m); prev_m = t; }
      if (do_r && (k <= 2 || done)) { xr -= mp.rate[tr]*(t - prev_r); prev_r = t; }
      if (done) { ++v; need_init = true; }
      else {
        if (k == 3 && do_l) xl += mp.log_rate[tl];      // mid of the left triple jumped
        if (k == 2) xm += mp.log_rate[tm];
        if (k == 1 && do_r) xr += mp.log_rate[tr];
        bits ^= 1 << k;
        iE += k == 0; iC += k == 1; iP += k == 2; iB += k == 3; iA += k == 4;
      }
    }
  }

  // ---- accept / reject (log_accept_rate, SingleSiteSampler.cpp:396-433, :520-533) ---
  if (!live) return active ? status : 0;
  const double llr = (xl + xr) - l_old - r_old;
  Philox rng0; rng0.init(seed, sweep, gsite, 0u);
  const double u = rng0.uniform();
  if (!(llr >= 0.0 || u < exp(llr))) return 0;
  tri_llh[s - 1] = xl; tri_llh[s] = xm; tri_llh[s + 1] = xr;
  for (int w = 0; w < c.W; ++w) {
    unsigned sb = 0;
    for (int k = 0; k < 8; ++k) {
      const int v = w*8 + k;
      if (v >= 1 && v < n_nodes) sb |= (unsigned)pr.pend[c.tree.parent[v]] << k;
    }
    state_rw[s*c.W + w] = (uint8_t)sb;
  }
  return 1;
}

}  // namespace epv
