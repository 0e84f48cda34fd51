// mh_mix.cuh -- the Metropolis-Hastings site update, third control-flow design (round 2):
// LOCKSTEP for the branches whose neighbours hold no jumps, PER-LANE FLAT LOOPS for the rest.
//
// What the two earlier designs showed (ncu, B200, 8-leaf tree, 0.05 jumps per row):
//   mh_site_flat  (mh_flat.cuh)  every lane walks ALL its (branch, interval) pairs with its own cursor:
//                 18.4 k warp instructions per warp, 18.9 lanes per instruction, 5.1 ms per pass;
//   mh_site_lock  (mh_lock.cuh)  branch loop warp-uniform, single-interval branches from a table:
//                 the cheap part got cheap, but the interval loop of the few lanes with neighbour jumps
//                 ran once per BRANCH (2.2 heavy iterations per branch at 2.8 lanes): 24.6 k warp
//                 instructions per warp at 13.2 lanes, 6.6 ms per pass.
// Here the heavy work is taken out of the tree order so that each lane can run it in ONE flat loop over
// its own few multi-interval branches (max over lanes of ~3 intervals instead of 14 x 2.2):
//   * the transfer matrix of a branch, M_v = P_1 P_2 ... P_k (2x2 stochastic: two numbers), and the
//     suffix products S_i = P_i ... P_k depend on the NEIGHBOURS' rows only, not on Felsenstein's q --
//     phase B computes them per lane, flat, before the tree is touched;
//   * upward pruning (phase U) and the draw of every branch's END state (phase D) are then uniform over
//     branches: P from the per-CTA table [node][4] (single interval) or M_v (per-lane array), no inner
//     loop, the branch index warp-uniform (tree tables, q of internal nodes in shared memory [slot][thread],
//     leaf q as one bit, end states as a bit mask);
//   * the states at the interval boundaries INSIDE a multi-interval branch are drawn afterwards (phase R,
//     per lane, flat, forward in time) from  P(x_{i+1} | x_i, end) ~ P_i(x_i, x_{i+1}) S_{i+1}(x_{i+1}, end)
//     -- the same joint law as drawing them root to leaf against the partial vectors (SingleSiteSampler.cpp
//     :180-221): prod_i P_i(x_i, x_{i+1}) / M_v(x_0, end) times the end-state law of phase D;
//   * jump times: queued intervals, warp-cooperative sampling (mh_flat.cuh), unchanged;
//   * the three triple log-likelihoods: closed form for branches whose five rows and proposal are empty
//     (lockstep), one flat five-way merge loop per lane over the other branches.
// Philox streams: (site, 1) one block per BRANCH for the end state and -- single-interval branches -- the
// no-jump draw; (site, 2^20 + v) one block per interval of multi-interval branch v; (site, 2 + k) the k-th
// queued interval; (site, 0) the acceptance draw.  Same law as mh_site_flat, not the same chain.
#pragma once
#include "mh_lock.cuh"

namespace epv {

template <int MAXN, int PB, int QCAP, bool QSMEM>
struct MixScratch {
  double q_local[2*MAXN];   // !QSMEM; also the generic fallback's q
  double stk[3*PB];         // per interval of the multi-interval branches: (h_i, S_{i+1}(0,0), S_{i+1}(1,1))
  double mv[2*PB];          // per multi-interval branch, in the order of phase B: M_v(0,0), M_v(1,1)
  double *q_s;              // this thread's column of the shared q array
  uint16_t *pn_s;           // ... of the shared jump counts per node
  uint16_t *poff_s;         // ... of the shared start offsets (into pj) per node
  QSlot *queue;
  int *excl;
};

template <typename M> __device__ __forceinline__ int top_bit(M x);
template <> __device__ __forceinline__ int top_bit<uint32_t>(uint32_t x) { return 31 - __clz((int)x); }
template <> __device__ __forceinline__ int top_bit<uint64_t>(uint64_t x) { return 63 - __clzll((long long)x); }
template <typename M> __device__ __forceinline__ int low_bit(M x);
template <> __device__ __forceinline__ int low_bit<uint32_t>(uint32_t x) { return __ffs((int)x) - 1; }
template <> __device__ __forceinline__ int low_bit<uint64_t>(uint64_t x) { return __ffsll((long long)x) - 1; }
__device__ __forceinline__ int pop_count(uint32_t x) { return __popc(x); }
__device__ __forceinline__ int pop_count(uint64_t x) { return __popcll(x); }

// collect_queue with per-node counts AND start offsets (the queue holds two ascending runs of branches:
// single-interval branches from phase D, then multi-interval ones from phase R; the items of one branch
// are contiguous).
template <typename M>
__device__ __forceinline__ int
collect_queue_mix(const ModelParams &mp, QSlot *queue, int &qn, uint64_t seed, uint32_t sweep, uint64_t gsite,
                  double *pj, int cap, uint16_t *pn_s, uint16_t *poff_s, M &propmask, int total) {
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
    const int add = total - before;
    if (add > 0) {
      if ((propmask >> v) & 1) {
        const int sum = add + (int)pn_s[v*LOCK_STRIDE];
        if (sum > 65535) return -1;
        pn_s[v*LOCK_STRIDE] = (uint16_t)sum;
      }
      else {
        if (before > 65535 || add > 65535) return -1;
        pn_s[v*LOCK_STRIDE] = (uint16_t)add; poff_s[v*LOCK_STRIDE] = (uint16_t)before;
        propmask |= (M)1 << v;
      }
    }
  }
  qn = 0;
  return total;
}

// Return values and the warp-synchronous calling convention are those of mh_site_flat.  On return pend
// holds the end state of every branch (bit v; bit 0 = root state), propmask the nodes whose new row holds
// jumps: S.pn_s[v] of them, starting at pj[S.poff_s[v]].
template <int MAXN, int PB, int QCAP, bool HALO, bool QSMEM>
__device__ __forceinline__ int
mh_site_mix(bool active, bool coop, const ModelParams &mp, const CtxTables &tb, const TreeTab *tt,
            const BranchTab *bt, const SiteCtx &c, uint8_t *state_rw, double *tri_llh, uint64_t seed,
            uint32_t sweep, uint64_t gsite, uint64_t n_global, MixScratch<MAXN, PB, QCAP, QSMEM> &S,
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

  // ---- 0. classification of the branches (lockstep) -------------------------------------------------
  M nbr = 0, llhm = 0, leafst = 0;          // neighbours hold jumps / any of the four rows does / leaf state
  for (int v = 1; v < n_nodes; ++v) {
    Row L, R; int own_n = 0, nA = 0, nE = 0;
    L.n = R.n = 0;
    if (ok) load_five<HALO>(c.pv, sw, v - 1, s, do_l, do_r, L, R, own_n, nA, nE);
    if ((L.n | R.n) != 0) nbr |= (M)1 << v;
    if ((L.n | R.n | nA | nE) != 0) llhm |= (M)1 << v;
    if (tt[v].leaf) leafst |= (M)(sw.bit(0, v) ^ (own_n & 1)) << v;
  }
  if (pop_count(nbr) > PB) { status = -3; ok = false; }

  // ---- B. multi-interval branches: suffix products and M_v, flat per lane, backward in time ----------
  int sp = 0;
  {
    M todo = ok ? nbr : 0;
    int v = todo ? top_bit<M>(todo) : 0, nm = 0;
    RevMerge m;
    double s00 = 1.0, s11 = 1.0;
    bool need_init = true;
    while (__any_sync(FULL, v >= 1)) {
      if (v < 1) continue;
      if (need_init) {
        Row L, R; int own_n;
        load_lmr<HALO>(c.pv, sw, v - 1, s, L, R, own_n);
        m.start(L, R, tt[v].len);
        s00 = 1.0; s11 = 1.0; need_init = false;
      }
      double len; int ctx; bool first;
      m.next(len, ctx, first);
      const double h = exp(-len*tb.s[ctx]);
      double P00, P11;
      stay_probs(tb, ctx, h, P00, P11);
      if (sp >= PB) { status = -3; ok = false; v = 0; continue; }
      S.stk[3*sp] = h; S.stk[3*sp + 1] = s00; S.stk[3*sp + 2] = s11; ++sp;
      const double n00 = P00*s00 + (1.0 - P00)*(1.0 - s11), n11 = (1.0 - P11)*(1.0 - s00) + P11*s11;
      s00 = n00; s11 = n11;
      if (first) {
        S.mv[2*nm] = s00; S.mv[2*nm + 1] = s11; ++nm;
        todo &= ~((M)1 << v);
        v = todo ? top_bit<M>(todo) : 0;
        need_init = true;
      }
    }
  }

  // ---- U. upward pruning, lockstep: one transfer matrix per branch -------------------------------------
  for (int v = 0; v < n_nodes; ++v) {
    const int slot = tt[v].qslot;
    if (slot >= 0) { qref(v, slot, 0) = 1.0; qref(v, slot, 1) = 1.0; }
  }
  for (int v = n_nodes - 1; v >= 1; --v) {
    const TreeTab t = tt[v];
    if (t.parent == 0) continue;                       // the root's q is never used (proposal identity)
    double p0, p1;
    if (t.leaf) { const int ls = (int)((leafst >> v) & 1); p0 = ls ? 0.0 : 1.0; p1 = ls ? 1.0 : 0.0; }
    else { p0 = qref(v, t.qslot, 0); p1 = qref(v, t.qslot, 1); }
    double P00, P11;
    if ((nbr >> v) & 1) { const int ord = pop_count((M)((nbr >> v) >> 1)); P00 = S.mv[2*ord]; P11 = S.mv[2*ord + 1]; }
    else { const int c4 = sw.bit(-1, v)*2 + sw.bit(1, v); P00 = bt[v].P00[c4]; P11 = bt[v].P11[c4]; }
    const double a = P00*p0 + (1.0 - P00)*p1, b = (1.0 - P11)*p0 + P11*p1;
    if (ok) {
      const int ps = tt[t.parent].qslot;
      qref(t.parent, ps, 0) *= a; qref(t.parent, ps, 1) *= b;
    }
  }

  // ---- D. end state of every branch, lockstep; single-interval branches are finished here ------------
  Philox rng; rng.init(seed, sweep, gsite, 1u);
  M pend = active ? (M)sw.bit(0, 1) : 0, propmask = 0;
  int total = 0, qn = 0, q_total = 0;
  for (int v = 1; v < n_nodes; ++v) {
    const TreeTab t = tt[v];
    const bool nontriv = (nbr >> v) & 1;
    const int prev = (int)((pend >> t.parent) & 1);
    const int c4 = sw.bit(-1, v)*2 + sw.bit(1, v);
    double P00, P11;
    if (nontriv) { const int ord = pop_count((M)((nbr >> v) >> 1)); P00 = S.mv[2*ord]; P11 = S.mv[2*ord + 1]; }
    else { P00 = bt[v].P00[c4]; P11 = bt[v].P11[c4]; }
    double nx0, nx1;
    if (t.leaf) { const int ls = (int)((leafst >> v) & 1); nx0 = ls ? 0.0 : 1.0; nx1 = ls ? 1.0 : 0.0; }
    else { nx0 = qref(v, t.qslot, 0); nx1 = qref(v, t.qslot, 1); }
    const double to0 = prev ? 1.0 - P11 : P00, to1 = prev ? P11 : 1.0 - P00;
    const double a0 = to0*nx0, a1 = to1*nx1;
    double u_end, u_jump;
    rng.uniform2(u_end, u_jump);
    const int e = u_end*(a0 + a1) > a0;                    // end state of the branch
    pend |= (M)e << v;
    if (ok && !nontriv) {
      // one interval: jumps iff the ends differ, or -- equal ends -- the "no jump at all" draw fails:
      // P(no jump | start == end) = exp(-r_prev len) / P(prev -> prev)
      bool need_jumps = e != prev;
      if (!need_jumps) need_jumps = !(u_jump*(prev ? P11 : P00) < (prev ? bt[v].e1[c4] : bt[v].e0[c4]));
      if (need_jumps) {
        if (qn == QCAP) total = collect_queue_mix(mp, S.queue, qn, seed, sweep, gsite, pj, pj_cap, S.pn_s, S.poff_s, propmask, total);
        if (total < 0) { status = total; ok = false; }
        else {
          QSlot it; it.a = 0.0; it.b = t.len; it.pad = 0;
          it.meta = (unsigned)v | (unsigned)ctx_of4(c4) << 8 | (unsigned)prev << 11 | (unsigned)e << 12 |
                    ((unsigned)(q_total++) & 0x7fffu) << 16;
          S.queue[qn++] = it;
        }
      }
    }
  }

  // ---- R. multi-interval branches: boundary states and queued intervals, flat per lane, forward -------
  {
    M todo = ok ? nbr : 0;
    int v = todo ? low_bit<M>(todo) : n_nodes, x = 0, e = 0;
    FwdMerge m;
    Philox rr; rr.init(seed, sweep, gsite, 0u);
    double passed = 0.0;
    bool need_init = true;
    while (__any_sync(FULL, v < n_nodes)) {
      if (v >= n_nodes) continue;
      if (need_init) {
        Row L, R; int own_n;
        load_lmr<HALO>(c.pv, sw, v - 1, s, L, R, own_n);
        m.start(L, R, tt[v].len);
        x = (int)((pend >> tt[v].parent) & 1); e = (int)((pend >> v) & 1);
        passed = 0.0;
        rr.init(seed, sweep, gsite, (1u << 20) | (unsigned)v);
        need_init = false;
      }
      double len; int ctx; bool last;
      m.next(len, ctx, last);
      --sp;
      const double h = S.stk[3*sp], s00 = S.stk[3*sp + 1], s11 = S.stk[3*sp + 2];
      double P00, P11;
      stay_probs(tb, ctx, h, P00, P11);
      double u_state, u_jump;
      rr.uniform2(u_state, u_jump);
      int y = e;
      if (!last) {
        const double w0 = (x ? 1.0 - P11 : P00)*(e ? 1.0 - s00 : s00), w1 = (x ? P11 : 1.0 - P00)*(e ? s11 : 1.0 - s11);
        y = u_state*(w0 + w1) > w0;
      }
      bool need_jumps = y != x;
      if (!need_jumps) {
        // compare u P(x -> x) with exp(-r_x len) through the alternating-series bounds of exp; exp itself
        // only in the gap (the decision is the one exp gives: the margin covers the rounding of the bounds)
        const double z = len*(x ? mp.rate[ctx | 2] : mp.rate[ctx]);
        const double lhs = u_jump*(x ? P11 : P00);
        const double lo = 1.0 + z*(-1.0 + z*(0.5 - z*(1.0/6.0)));
        const double z2 = z*z;
        if (lhs < lo - 1e-15) need_jumps = false;
        else if (lhs > lo + z2*z2*(1.0/24.0) + 1e-15) need_jumps = true;
        else need_jumps = !(lhs < exp(-z));
      }
      if (need_jumps) {
        if (qn == QCAP) total = collect_queue_mix(mp, S.queue, qn, seed, sweep, gsite, pj, pj_cap, S.pn_s, S.poff_s, propmask, total);
        if (total < 0) { status = total; ok = false; v = n_nodes; continue; }
        QSlot it; it.a = passed; it.b = len; it.pad = 0;
        it.meta = (unsigned)v | (unsigned)ctx << 8 | (unsigned)x << 11 | (unsigned)y << 12 |
                  ((unsigned)(q_total++) & 0x7fffu) << 16;
        S.queue[qn++] = it;
      }
      passed += len; x = y;
      if (last) {
        todo &= ~((M)1 << v);
        v = todo ? low_bit<M>(todo) : n_nodes;
        need_init = true;
      }
    }
  }

  // ---- S. jump times of the queued intervals ---------------------------------------------------------
  if (!ok) qn = 0;
  if (coop) coop_sample<QCAP>(mp, S.queue, S.excl, qn, seed, sweep, gsite);
  if (ok) {
    total = collect_queue_mix(mp, S.queue, qn, seed, sweep, gsite, pj, pj_cap, S.pn_s, S.poff_s, propmask, total);
    if (total < 0) { status = total; ok = false; }
  }
  __syncwarp(FULL);
  total_out = total; pend_out = pend; propmask_out = propmask;

  // ---- L. the three triple log-likelihoods ----------------------------------------------------------------
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
  const M busy = live ? (M)((llhm | propmask) & ~(M)1) : 0;      // branches that need the merge
  // branches whose five rows and proposal are empty: the single `done' event of the merge, in closed form
  for (int v = 1; v < n_nodes; ++v) {
    const TreeTab t = tt[v];
    if (live && !((busy >> v) & 1)) {
      const int Pinit = (int)((pend >> t.parent) & 1);
      const int bits = (dl ? sw.bit(-2, v) : 0) << 4 | sw.bit(-1, v) << 3 | Pinit << 2 | sw.bit(1, v) << 1 |
                       (dr ? sw.bit(2, v) : 0);
      if (dl) xl -= mp.rate[bits >> 2]*t.len;
      xm -= mp.rate[(bits >> 1) & 7]*t.len;
      if (dr) xr -= mp.rate[bits & 7]*t.len;
    }
  }
  {
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    M todo = busy;
    int v = todo ? low_bit<M>(todo) : n_nodes;
    Row A, B, P, C, E;
    A.n = B.n = P.n = C.n = E.n = 0; A.ptr = B.ptr = P.ptr = C.ptr = E.ptr = pj;
    int iA = 0, iB = 0, iP = 0, iC = 0, iE = 0, bits = 0;   // bits: A<<4 | B<<3 | P<<2 | C<<1 | E
    double prev_l = 0.0, prev_m = 0.0, prev_r = 0.0, tot = 0.0;
    bool need_init = true;
    while (__any_sync(FULL, v < n_nodes)) {
      if (v >= n_nodes) continue;
      if (need_init) {
        const int bi = v - 1;
        row_data<HALO>(c.pv, bi, s - 1, B.ptr, B.n);
        row_data<HALO>(c.pv, bi, s + 1, C.ptr, C.n);
        if (dl) row_data<HALO>(c.pv, bi, s - 2, A.ptr, A.n); else { A.n = 0; A.ptr = B.ptr; }
        if (dr) row_data<HALO>(c.pv, bi, s + 2, E.ptr, E.n); else { E.n = 0; E.ptr = C.ptr; }
        const bool has = (propmask >> v) & 1;
        P.n = has ? (int)S.pn_s[v*LOCK_STRIDE] : 0;
        P.ptr = pj + (has ? (int)S.poff_s[v*LOCK_STRIDE] : 0);
        iA = iB = iP = iC = iE = 0;
        bits = (dl ? sw.bit(-2, v) : 0) << 4 | sw.bit(-1, v) << 3 | (int)((pend >> tt[v].parent) & 1) << 2 |
               sw.bit(1, v) << 1 | (dr ? sw.bit(2, v) : 0);
        prev_l = prev_m = prev_r = 0.0;
        tot = tt[v].len;
        need_init = false;
      }
      // next event: smallest time, ties to the higher site index (the reference's
      // right-before-mid-before-left rule, Path.cpp:275-281, holds for all three triples)
      double tmin = iE < E.n ? E.ptr[iE] : INF; int k = 0;
      { const double z = iC < C.n ? C.ptr[iC] : INF; if (z < tmin) { tmin = z; k = 1; } }
      { const double z = iP < P.n ? P.ptr[iP] : INF; if (z < tmin) { tmin = z; k = 2; } }
      { const double z = iB < B.n ? B.ptr[iB] : INF; if (z < tmin) { tmin = z; k = 3; } }
      { const double z = iA < A.n ? A.ptr[iA] : INF; if (z < tmin) { tmin = z; k = 4; } }
      const bool done = !(tmin < INF);
      if (done) tmin = tot;
      const int tl = bits >> 2, tm = (bits >> 1) & 7, tr = bits & 7;
      if (dl && (k >= 2 || done)) { xl -= mp.rate[tl]*(tmin - prev_l); prev_l = tmin; }
      if ((k >= 1 && k <= 3) || done) { xm -= mp.rate[tm]*(tmin - prev_m); prev_m = tmin; }
      if (dr && (k <= 2 || done)) { xr -= mp.rate[tr]*(tmin - prev_r); prev_r = tmin; }
      if (done) {
        todo &= ~((M)1 << v);
        v = todo ? low_bit<M>(todo) : n_nodes;
        need_init = true;
      }
      else {
        if (k == 3 && dl) xl += mp.log_rate[tl];      // mid of the left triple jumped
        if (k == 2) xm += mp.log_rate[tm];
        if (k == 1 && dr) xr += mp.log_rate[tr];
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
