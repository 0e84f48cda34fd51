// mh_lock.cuh -- the Metropolis-Hastings site update with the BRANCH loop in lockstep.
//
// Same mathematics, same Philox streams and the same per-site order of operations as
// mh_site_flat (mh_flat.cuh; SingleSiteSampler.cpp:482-536), so a site's chain is unchanged;
// different control flow.  mh_site_flat gives every lane its own (branch, interval) cursor.
// Here the 32 sites of a warp walk the tree TOGETHER (the branch index v is warp-uniform):
//
//   * ~90 % of (site, branch) pairs have neighbours without jumps on the branch: ONE interval
//     of length exactly branch[v] in one of four contexts.  Its transition probabilities
//     P00 / P11 and the two no-jump probabilities exp(-rate len) come from a per-CTA shared
//     table [node][4] filled once per launch with the very expressions the interval loop uses
//     (same bits) -- no exp, no offset -> pointer -> jump-time chain, no interval stack;
//   * the first pass classifies every branch of the site (neighbours hold jumps? any of the
//     four rows s-2, s-1, s+1, s+2 holds jumps?) into bit masks kept in registers; the later
//     passes touch the CSR only for the branches that need it;
//   * a branch whose five rows and whose proposal are all empty contributes
//     -(rate[tl] + rate[tm] + rate[tr]) * branch[v] to the three triple log-likelihoods in
//     closed form (the one `done' event of the five-way merge);
//   * lanes that do have neighbour jumps run the interval loop of the branch under
//     `while (__any_sync)` -- iteration 0 is shared with the single-interval lanes;
//   * tree tables (parent, leaf flag, length, q slot) are warp-uniform shared-memory reads;
//     Felsenstein q is kept for INTERNAL nodes only (a leaf's q is an indicator: one bit in a
//     register) in shared memory laid out [slot][thread] (bank = thread), end states and the
//     "this branch got jumps" flags are 64-bit masks in registers.
//
// Trees of up to 64 nodes (one 64-bit mask per quantity); wider trees keep mh_site_flat.
#pragma once
#include "mh_flat.cuh"

namespace epv {

struct TreeTab {            // per node, shared memory, read warp-uniformly
  double len;               // branch length above the node (= tot_time of its rows)
  int parent;
  short leaf;               // subtree size 1
  short qslot;              // slot of the node's q in shared memory (internal non-root nodes), else -1
};

struct BranchTab {          // per node x context (ctx4 = l*2 + r): the single-interval case
  double P00[4], P11[4];    // transition probabilities over the whole branch
  double e0[4], e1[4];      // exp(-rate[ctx] len), exp(-rate[ctx|2] len): P(no jump at all)
};

__device__ __forceinline__ int ctx_of4(int ctx4) { return ((ctx4 & 2) << 1) | (ctx4 & 1); }

// The one expression for the 2x2 transition probabilities of an interval with exp term h
// (ContinuousTimeMarkovModel.cpp:115-125 with pi0 = r0/s, pi1 = r1/s): used by the table and
// by the interval loops, so both give the same bits.
__device__ __forceinline__ void
stay_probs(const CtxTables &tb, int ctx, double h, double &P00, double &P11) {
  P00 = tb.pi1[ctx] + tb.pi0[ctx]*h;
  P11 = tb.pi0[ctx] + tb.pi1[ctx]*h;
}

// Fills the per-CTA tables (all threads of the CTA; the caller synchronises before and after).
__device__ __forceinline__ void
fill_lock_tables(const ModelParams &mp, const CtxTables &tb, const TreeDev &tree, TreeTab *tt, BranchTab *bt,
                 int t, int n_threads) {
  const int n = tree.n_nodes;
  for (int i = t; i < 4*n; i += n_threads) {
    const int v = i >> 2, c4 = i & 3, ctx = ctx_of4(c4);
    const double len = v ? tree.branch[v] : 0.0;
    const double h = exp(-len*tb.s[ctx]);
    double P00, P11;
    stay_probs(tb, ctx, h, P00, P11);
    bt[v].P00[c4] = P00; bt[v].P11[c4] = P11;
    bt[v].e0[c4] = exp(-(len*mp.rate[ctx]));
    bt[v].e1[c4] = exp(-(len*mp.rate[ctx | 2]));
  }
  if (t == 0) {
    int slots = 0;
    for (int v = 0; v < n; ++v) {
      tt[v].len = v ? tree.branch[v] : 0.0;
      tt[v].parent = tree.parent[v];
      tt[v].leaf = tree.subtree[v] == 1;
      tt[v].qslot = (short)((v > 0 && tree.subtree[v] > 1) ? slots++ : -1);
    }
  }
}

// Number of q slots a tree needs (host side: sizes the shared memory).
inline int lock_q_slots(int n_nodes, const int *subtree) {
  int slots = 0;
  for (int v = 1; v < n_nodes; ++v) slots += subtree[v] > 1;
  return slots;
}

// Per-thread scratch of the lockstep routine.  q: shared memory [2*slot + k][thread] when QSMEM,
// else a local array indexed by node like mh_site_flat's.
template <int MAXN, int PB, int QCAP, bool QSMEM>
struct LockScratch {
  double q_local[2*MAXN];  // !QSMEM; also the generic fallback's q (unused arrays are dropped by the compiler)
  double hb[MAXN];          // exp term of the FIRST interval of multi-interval branches only
  double stk[3*PB];         // (p0, p1, h) of the later intervals of multi-interval branches
  double *q_s;              // this thread's column of the shared q array (stride MH_THREADS_ doubles)
  uint16_t *pn_s;           // this thread's column of the shared jump counts per node (stride MH_THREADS_)
  QSlot *queue;
  int *excl;
};

constexpr int LOCK_STRIDE = 128;     // == MH_THREADS (kernels.h); columns of the shared per-thread arrays
template <int MAXN> struct LockMask { typedef uint64_t type; };
template <> struct LockMask<8> { typedef uint32_t type; };
template <> struct LockMask<16> { typedef uint32_t type; };
template <> struct LockMask<32> { typedef uint32_t type; };

// Init-state bits of sites s-2..s+2 in the narrowest word that holds the tree's nodes.
template <typename M>
struct StateWinM {
  M w[5];
  __device__ __forceinline__ void
  load(const uint8_t *state, int W, uint64_t s, uint64_t n_sites) {
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      w[k] = 0;
      const int64_t site = (int64_t)s + k - 2;
      if (site >= 0 && site < (int64_t)n_sites) w[k] = (M)state_word(state, W, (uint64_t)site, 0);
    }
  }
  __device__ __forceinline__ int bit(int d, int node) const { return (int)((w[d + 2] >> node) & 1u); }
};

// Counts of the rows s-2 .. s+2 on branch bi and the two neighbour rows' data.
template <bool HALO, typename SW>
__device__ __forceinline__ void
load_five(const PathView &pv, const SW &sw, int bi, uint64_t site, bool do_l, bool do_r,
          Row &L, Row &R, int &own_n, int &nA, int &nE) {
  L.init = sw.bit(-1, bi + 1);
  R.init = sw.bit(1, bi + 1);
  if (!HALO) {
    const uint32_t *o = pv.off + (uint64_t)bi*pv.off_pitch + (site - 1 - pv.own_lo);
    const uint32_t o0 = o[0], o1 = o[1], o2 = o[2], o3 = o[3];
    const uint32_t oA = do_l ? o[-1] : o0, oE = do_r ? o[4] : o3;
    const double *base = pv.jmp + pv.jbase[bi];
    L.ptr = base + o0; L.n = (int)(o1 - o0);
    own_n = (int)(o2 - o1);
    R.ptr = base + o2; R.n = (int)(o3 - o2);
    nA = (int)(o0 - oA); nE = (int)(oE - o3);
  }
  else {
    const double *dummy;
    row_data<true>(pv, bi, site - 1, L.ptr, L.n);
    row_data<true>(pv, bi, site + 1, R.ptr, R.n);
    row_data<true>(pv, bi, site, dummy, own_n);
    nA = nE = 0;
    if (do_l) row_data<true>(pv, bi, site - 2, dummy, nA);
    if (do_r) row_data<true>(pv, bi, site + 2, dummy, nE);
  }
}

// collect_queue of mh_flat.cuh with the jump counts per node in shared memory and a mask of the
// nodes that hold any (no zero-initialisation pass over the counts).
template <typename M>
__device__ __forceinline__ int
collect_queue_lock(const ModelParams &mp, QSlot *queue, int &qn, uint64_t seed, uint32_t sweep, uint64_t gsite,
                   double *pj, int cap, uint16_t *pn_s, M &propmask, int total) {
  for (int k = 0; k < qn; ++k) {
    const QSlot it = queue[k];
    const int v = it.meta & 0xff;
    const int before = total;
    if (it.meta & Q_DONE) {
      const int cnt = (it.meta >> 16) & 3;
      if (total + cnt > cap) return -1;
      pj[total] = it.a;
      if (cnt > 1) pj[total + 1] = it.b;
      total += cnt;
    }
    else {
      total = sample_slot(mp, it, seed, sweep, gsite, pj, total, cap);
      if (total < 0) return total;
    }
    const int add = total - before + (((propmask >> v) & 1) ? (int)pn_s[v*LOCK_STRIDE] : 0);
    if (add > 65535) return -1;
    if (add > 0) { pn_s[v*LOCK_STRIDE] = (uint16_t)add; propmask |= (M)1 << v; }
  }
  qn = 0;
  return total;
}

// Return values and the warp-synchronous calling convention are those of mh_site_flat.  On return
// pend holds the end state of every branch (bit v; bit 0 = root state), propmask the nodes whose new
// row holds jumps (their counts are in S.pn_s), pj[0 .. total_out) the new jump times in preorder.
template <int MAXN, int PB, int QCAP, bool HALO, bool QSMEM>
__device__ __forceinline__ int
mh_site_lock(bool active, bool coop, const ModelParams &mp, const CtxTables &tb, const TreeTab *tt,
             const BranchTab *bt, const SiteCtx &c, uint8_t *state_rw, double *tri_llh, uint64_t seed,
             uint32_t sweep, uint64_t gsite, uint64_t n_global, LockScratch<MAXN, PB, QCAP, QSMEM> &S,
             double *pj, int pj_cap, typename LockMask<MAXN>::type &pend_out,
             typename LockMask<MAXN>::type &propmask_out, int &total_out) {
  static_assert(MAXN <= 64, "one 64-bit mask per quantity");
  typedef typename LockMask<MAXN>::type M;
  constexpr unsigned FULL = 0xffffffffu;
  const int n_nodes = c.tree.n_nodes;
  const uint64_t s = c.site;
  int status = 1;
  bool ok = active;
  const bool do_l = gsite > 1, do_r = gsite + 2 < n_global;
  StateWinM<M> sw;
  if (active) sw.load(c.state, c.W, s, c.pv.n_local);
  else { for (int k = 0; k < 5; ++k) sw.w[k] = 0; }
  if (!HALO && active) asm volatile("prefetch.global.L2 [%0];" :: "l"(tri_llh + s - 1));

  auto qref = [&](int v, int slot, int k) -> double & {
    return QSMEM ? S.q_s[(2*slot + k)*LOCK_STRIDE] : S.q_local[2*v + k];
  };

  // ---- 1. upward pruning; classification of the branches ------------------------------
  M nbr = 0, llhm = 0, leafst = 0;
  int sp = 0;
  for (int v = 0; v < n_nodes; ++v) {
    const int slot = tt[v].qslot;
    if (slot >= 0) { qref(v, slot, 0) = 1.0; qref(v, slot, 1) = 1.0; }
  }
  for (int v = n_nodes - 1; v >= 1; --v) {
    const TreeTab t = tt[v];
    Row L, R; int own_n = 0, nA = 0, nE = 0;
    L.n = R.n = 0; L.init = R.init = 0; L.ptr = R.ptr = pj;
    if (ok) load_five<HALO>(c.pv, sw, v - 1, s, do_l, do_r, L, R, own_n, nA, nE);
    const bool nontriv = (L.n | R.n) != 0;
    if (nontriv) nbr |= (M)1 << v;
    if ((L.n | R.n | nA | nE) != 0) llhm |= (M)1 << v;
    double p0, p1;
    if (t.leaf) {
      const int ls = sw.bit(0, v) ^ (own_n & 1);
      leafst |= (M)ls << v;
      p0 = ls ? 0.0 : 1.0; p1 = ls ? 1.0 : 0.0;
    }
    else { p0 = qref(v, t.qslot, 0); p1 = qref(v, t.qslot, 1); }
    RevMerge m;
    if (nontriv) m.start(L, R, t.len);
    bool more = ok;
    do {
      if (more) {
        double P00, P11, h = 0.0;
        bool first = true;
        if (!nontriv) { const int c4 = L.init*2 + R.init; P00 = bt[v].P00[c4]; P11 = bt[v].P11[c4]; }
        else {
          double len; int ctx;
          m.next(len, ctx, first);
          h = exp(-len*tb.s[ctx]);
          stay_probs(tb, ctx, h, P00, P11);
        }
        const double a = P00*p0 + (1.0 - P00)*p1, b = (1.0 - P11)*p0 + P11*p1;
        p0 = a; p1 = b;
        if (first) { if (nontriv) S.hb[v] = h; more = false; }
        else if (sp >= PB) { status = -3; ok = false; more = false; }
        else { S.stk[3*sp] = p0; S.stk[3*sp + 1] = p1; S.stk[3*sp + 2] = h; ++sp; }
      }
    } while (__any_sync(FULL, more));
    if (t.parent > 0 && ok) {
      const int ps = tt[t.parent].qslot;
      qref(t.parent, ps, 0) *= p0; qref(t.parent, ps, 1) *= p1;
    }
  }

  // ---- 2. downward pass -----------------------------------------------------------------
  Philox rng; rng.init(seed, sweep, gsite, 1u);
  M pend = active ? (M)sw.bit(0, 1) : 0, propmask = 0;
  int total = 0, qn = 0, q_total = 0;
  for (int v = 1; v < n_nodes; ++v) {
    const TreeTab t = tt[v];
    const bool nontriv = (nbr >> v) & 1;
    int prev = (int)((pend >> t.parent) & 1), i = 0;
    double passed = 0.0, h_saved = 0.0;
    FwdMerge m;
    const int c4 = sw.bit(-1, v)*2 + sw.bit(1, v);
    if (ok && nontriv) {
      Row L, R; int own_n;
      load_lmr<HALO>(c.pv, sw, v - 1, s, L, R, own_n);
      m.start(L, R, t.len);
    }
    bool more = ok;
    do {
      if (more) {
        double len, P00, P11; int ctx; bool last = true;
        if (!nontriv) { len = t.len; ctx = ctx_of4(c4); P00 = bt[v].P00[c4]; P11 = bt[v].P11[c4]; }
        else {
          m.next(len, ctx, last);
          const double h = i == 0 ? S.hb[v] : h_saved;
          stay_probs(tb, ctx, h, P00, P11);
        }
        double nx0, nx1;
        if (last) {
          if (t.leaf) { const int ls = (int)((leafst >> v) & 1); nx0 = ls ? 0.0 : 1.0; nx1 = ls ? 1.0 : 0.0; }
          else { nx0 = qref(v, t.qslot, 0); nx1 = qref(v, t.qslot, 1); }
        }
        else { --sp; nx0 = S.stk[3*sp]; nx1 = S.stk[3*sp + 1]; h_saved = S.stk[3*sp + 2]; }
        const double to0 = prev ? 1.0 - P11 : P00, to1 = prev ? P11 : 1.0 - P00;
        const double a0 = to0*nx0, a1 = to1*nx1;
        double u_end, u_jump;
        rng.uniform2(u_end, u_jump);
        const int e = u_end*(a0 + a1) > a0;                  // end state of the interval
        bool need_jumps = e != prev;
        if (!need_jumps) {
          // P(no jump | start == end) = exp(-r_prev len) / P(prev -> prev): compare u P(prev -> prev)
          // with exp(-x), x = r_prev len
          const double lhs = u_jump*(prev ? P11 : P00);
          if (!nontriv) need_jumps = !(lhs < (prev ? bt[v].e1[c4] : bt[v].e0[c4]));
          else {
            // (alternating-series bounds of exp(-x); exp itself only in the gap -- the decision is the
            // one exp gives: the margin covers the rounding of the bounds)
            const double x = len*(prev ? mp.rate[ctx | 2] : mp.rate[ctx]);
            const double lo = 1.0 + x*(-1.0 + x*(0.5 - x*(1.0/6.0)));
            const double x2 = x*x;
            if (lhs < lo - 1e-15) need_jumps = false;
            else if (lhs > lo + x2*x2*(1.0/24.0) + 1e-15) need_jumps = true;
            else need_jumps = !(lhs < exp(-x));
          }
        }
        if (need_jumps) {
          if (qn == QCAP) total = collect_queue_lock(mp, S.queue, qn, seed, sweep, gsite, pj, pj_cap, S.pn_s, propmask, total);
          if (total < 0) { status = total; ok = false; more = false; }
          else {
            QSlot it; it.a = passed; it.b = len; it.pad = 0;
            it.meta = (unsigned)v | (unsigned)ctx << 8 | (unsigned)prev << 11 | (unsigned)e << 12 |
                      ((unsigned)(q_total++) & 0x7fffu) << 16;
            S.queue[qn++] = it;
          }
        }
        if (more) {
          passed += len; prev = e; ++i;
          if (last) more = false;
        }
      }
    } while (__any_sync(FULL, more));
    pend |= (M)prev << v;
  }
  // ---- 3. jump times of the queued intervals ---------------------------------------
  if (status <= 0) qn = 0;
  if (coop) coop_sample<QCAP>(mp, S.queue, S.excl, qn, seed, sweep, gsite);
  if (ok) {
    total = collect_queue_lock(mp, S.queue, qn, seed, sweep, gsite, pj, pj_cap, S.pn_s, propmask, total);
    if (total < 0) { status = total; ok = false; }
  }
  __syncwarp(FULL);
  total_out = total; pend_out = pend; propmask_out = propmask;

  // ---- 4. the three triple log-likelihoods ---------------------------------------------
  const bool live = ok;
  const bool dl = live && do_l, dr = live && do_r;
  const double l_old = live ? tri_llh[s - 1] : 0.0, r_old = live ? tri_llh[s + 1] : 0.0;
  double xl = 0.0, xm = 0.0, xr = 0.0;
  if (live) {
    const int sm = (int)(pend & 1);
    const int sl1 = sw.bit(-1, 1), sr1 = sw.bit(1, 1);
    xm = mp.T[sl1*2 + sm]*mp.T[sm*2 + sr1];
    xl = dl ? mp.T[sw.bit(-2, 1)*2 + sl1]*mp.T[sl1*2 + sm] : l_old;
    xr = dr ? mp.T[sm*2 + sr1]*mp.T[sr1*2 + sw.bit(2, 1)] : r_old;
  }
  {
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    int poff = 0;
    for (int v = 1; v < n_nodes; ++v) {
      const TreeTab t = tt[v];
      const double tot = t.len;
      const int Pn = ((propmask >> v) & 1) ? (int)S.pn_s[v*LOCK_STRIDE] : 0;
      const int Pinit = (int)((pend >> t.parent) & 1);
      int bits = (dl ? sw.bit(-2, v) : 0) << 4 | sw.bit(-1, v) << 3 | Pinit << 2 | sw.bit(1, v) << 1 |
                 (dr ? sw.bit(2, v) : 0);                 // A<<4 | B<<3 | P<<2 | C<<1 | E
      const bool merge = live && (((llhm >> v) & 1) || Pn != 0);
      if (live && !merge) {
        // all five rows and the proposal are empty: the single `done' event of the merge
        const int tl = bits >> 2, tm = (bits >> 1) & 7, tr = bits & 7;
        if (dl) xl -= mp.rate[tl]*(tot - 0.0);
        xm -= mp.rate[tm]*(tot - 0.0);
        if (dr) xr -= mp.rate[tr]*(tot - 0.0);
      }
      bool more = merge;
      if (__any_sync(FULL, more)) {
        Row A, B, P, C, E;
        A.n = B.n = P.n = C.n = E.n = 0; A.ptr = B.ptr = P.ptr = C.ptr = E.ptr = pj;
        int iA = 0, iB = 0, iP = 0, iC = 0, iE = 0;
        double prev_l = 0.0, prev_m = 0.0, prev_r = 0.0;
        if (more) {
          const int bi = v - 1;
          row_data<HALO>(c.pv, bi, s - 1, B.ptr, B.n);
          row_data<HALO>(c.pv, bi, s + 1, C.ptr, C.n);
          if (dl) row_data<HALO>(c.pv, bi, s - 2, A.ptr, A.n); else { A.n = 0; A.ptr = B.ptr; }
          if (dr) row_data<HALO>(c.pv, bi, s + 2, E.ptr, E.n); else { E.n = 0; E.ptr = C.ptr; }
          P.ptr = pj + poff; P.n = Pn;
        }
        while (__any_sync(FULL, more)) {
          if (!more) continue;
          // next event: smallest time, ties to the higher site index (the reference's
          // right-before-mid-before-left rule, Path.cpp:275-281, holds for all three triples)
          double tmin = iE < E.n ? E.ptr[iE] : INF; int k = 0;
          { const double x = iC < C.n ? C.ptr[iC] : INF; if (x < tmin) { tmin = x; k = 1; } }
          { const double x = iP < P.n ? P.ptr[iP] : INF; if (x < tmin) { tmin = x; k = 2; } }
          { const double x = iB < B.n ? B.ptr[iB] : INF; if (x < tmin) { tmin = x; k = 3; } }
          { const double x = iA < A.n ? A.ptr[iA] : INF; if (x < tmin) { tmin = x; k = 4; } }
          const bool done = !(tmin < INF);
          if (done) tmin = tot;
          const int tl = bits >> 2, tm = (bits >> 1) & 7, tr = bits & 7;
          if (dl && (k >= 2 || done)) { xl -= mp.rate[tl]*(tmin - prev_l); prev_l = tmin; }
          if ((k >= 1 && k <= 3) || done) { xm -= mp.rate[tm]*(tmin - prev_m); prev_m = tmin; }
          if (dr && (k <= 2 || done)) { xr -= mp.rate[tr]*(tmin - prev_r); prev_r = tmin; }
          if (done) more = false;
          else {
            if (k == 3 && dl) xl += mp.log_rate[tl];      // mid of the left triple jumped
            if (k == 2) xm += mp.log_rate[tm];
            if (k == 1 && dr) xr += mp.log_rate[tr];
            bits ^= 1 << k;
            iE += k == 0; iC += k == 1; iP += k == 2; iB += k == 3; iA += k == 4;
          }
        }
      }
      poff += Pn;
    }
  }

  // ---- accept / reject (log_accept_rate, SingleSiteSampler.cpp:396-433, :520-533) ---
  if (!live) return active ? status : 0;
  const double llr = (xl + xr) - l_old - r_old;
  Philox rng0; rng0.init(seed, sweep, gsite, 0u);
  const double u = rng0.uniform();
  if (!(llr >= 0.0 || u < exp(llr))) return 0;
  tri_llh[s - 1] = xl; tri_llh[s] = xm; tri_llh[s + 1] = xr;
  uint64_t sb = 0;                                    // new init state of node v = end state of its parent
  for (int v = 1; v < n_nodes; ++v) sb |= (uint64_t)((pend >> tt[v].parent) & 1) << v;
  uint8_t *dst = state_rw + s*c.W;
  switch (c.W) {
  case 1: *dst = (uint8_t)sb; break;
  case 2: *reinterpret_cast<uint16_t *>(dst) = (uint16_t)sb; break;
  case 4: *reinterpret_cast<uint32_t *>(dst) = (uint32_t)sb; break;
  default: *reinterpret_cast<uint64_t *>(dst) = sb; break;
  }
  return 1;
}

}  // namespace epv
