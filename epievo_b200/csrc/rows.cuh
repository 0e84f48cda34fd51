// rows.cuh -- row accessors, neighbour-interval merges, 2-state CTMC closed forms and
// the triple log-likelihood, shared by every kernel of the path.
#pragma once
#include "common.cuh"

namespace epv {

// One site's history on one branch: init state + sorted jump times (Path.hpp:33-57).
// ptr may point to global memory (stored rows) or to a thread's proposal buffer.
struct Row {
  const double *ptr;
  int n;
  int init;
};

__device__ __forceinline__ Row
load_row(const PathView &pv, const uint8_t *state, int W, int bi, uint64_t site) {
  Row r;
  r.init = state_bit(state, W, site, bi + 1);
  if (site >= pv.own_lo && site < pv.own_hi) {
    const uint32_t *o = pv.off + (uint64_t)bi*pv.off_pitch + (site - pv.own_lo);
    const uint32_t a = o[0], b = o[1];
    r.ptr = pv.jmp + pv.jbase[bi] + a;
    r.n = (int)(b - a);
  }
  else {
    const int slot = site < pv.own_lo ? (int)site : HALO_W + (int)(site - pv.own_hi);
    r.n = pv.hcnt[slot*pv.n_br + bi];
    r.ptr = pv.hjmp + (size_t)(slot*pv.n_br + bi)*pv.halo_k;
  }
  return r;
}

// 2x2 transition probabilities of a 2-state CTMC over time t with leaving rates
// r0 (0->1) and r1 (1->0): P00=(r0 h + r1)/(r0+r1), P11=(r0 + r1 h)/(r0+r1),
// h = exp(-t (r0+r1))  (ContinuousTimeMarkovModel.cpp:115-125, :142-161).
struct P2 { double p00, p01, p10, p11, h; };

__device__ __forceinline__ P2
trans_prob(double r0, double r1, double t) {
  P2 P;
  const double s = r0 + r1;
  P.h = exp(-t*s);
  P.p00 = (r0*P.h + r1)/s;
  P.p01 = 1.0 - P.p00;
  P.p11 = (r0 + r1*P.h)/s;
  P.p10 = 1.0 - P.p11;
  return P;
}

// Forward two-way merge of the left and right neighbours' jumps on one branch
// (collect_segment_info, Segment.cpp:35-79).  Yields intervals in time order; a left
// jump flips bit 4 of the context, a right jump bit 1; ties go to the right list
// first (strict '<' at Segment.cpp:47).  ctx = l*4 + r (mid bit clear) = trip0.
struct FwdMerge {
  const double *l, *r;
  int i, j, nl, nr, ctx;
  double prev, tot;
  __device__ __forceinline__ void
  start(const Row &L, const Row &R, double tot_time) {
    l = L.ptr; r = R.ptr; nl = L.n; nr = R.n; i = j = 0;
    ctx = L.init*4 + R.init; prev = 0.0; tot = tot_time;
  }
  __device__ __forceinline__ int n_intervals() const { return nl + nr + 1; }
  // Returns the next interval (len, context); last==true on the final one.
  __device__ __forceinline__ void
  next(double &len, int &c, bool &last) {
    c = ctx;
    if (i < nl && (j >= nr || l[i] < r[j])) {
      len = l[i] - prev; prev = l[i]; ++i; ctx ^= 4; last = false;
    }
    else if (j < nr) {
      len = r[j] - prev; prev = r[j]; ++j; ctx ^= 1; last = false;
    }
    else { len = tot - prev; last = true; }
  }
};

// The same intervals in reverse time order (for the upward recursion).
struct RevMerge {
  const double *l, *r;
  int i, j, ctx;
  double end;
  __device__ __forceinline__ void
  start(const Row &L, const Row &R, double tot_time) {
    l = L.ptr; r = R.ptr; i = L.n; j = R.n;
    ctx = (L.init ^ (L.n & 1))*4 + (R.init ^ (R.n & 1));
    end = tot_time;
  }
  __device__ __forceinline__ bool done() const { return i < 0; }
  // Returns interval (len, context) ending at the current 'end'; first==true when it
  // is the branch's first interval (starts at time 0).
  __device__ __forceinline__ void
  next(double &len, int &c, bool &first) {
    c = ctx;
    if (i > 0 && (j == 0 || !(l[i-1] < r[j-1]))) {
      --i; len = end - l[i]; end = l[i]; ctx ^= 4; first = false;
    }
    else if (j > 0) {
      --j; len = end - r[j]; end = r[j]; ctx ^= 1; first = false;
    }
    else { len = end; first = true; i = -1; }
  }
};

// Complete-data log-likelihood contribution of one branch for the triple (L, M, R):
// sum over events of log(rate[ctx]) at mid jumps minus rate[ctx]*duration
// (add_sufficient_statistics, Path.cpp:267-301, folded with log_likelihood,
// SingleSiteSampler.cpp:342-352).  Tie-breaks as the reference: left only if strictly
// smallest, then mid if strictly before right, else right.
__device__ __forceinline__ double
branch_llh(const ModelParams &mp, const Row &L, const Row &M, const Row &R,
           double tot_time) {
  int trip = L.init*4 + M.init*2 + R.init;
  if ((L.n | M.n | R.n) == 0) return -mp.rate[trip]*tot_time;
  int i = 0, j = 0, k = 0;
  double prev = 0.0, x = 0.0;
  const double INF = __longlong_as_double(0x7ff0000000000000LL);
  for (;;) {
    const double tl = i < L.n ? L.ptr[i] : INF;
    const double tm = j < M.n ? M.ptr[j] : INF;
    const double tr = k < R.n ? R.ptr[k] : INF;
    if (i >= L.n && j >= M.n && k >= R.n) break;
    if (tl < fmin(tm, tr)) {
      x -= mp.rate[trip]*(tl - prev); prev = tl; trip ^= 4; ++i;
    }
    else if (tm < tr) {
      x -= mp.rate[trip]*(tm - prev); x += mp.log_rate[trip];
      prev = tm; trip ^= 2; ++j;
    }
    else {
      x -= mp.rate[trip]*(tr - prev); prev = tr; trip ^= 1; ++k;
    }
  }
  x -= mp.rate[trip]*(tot_time - prev);
  return x;
}

}  // namespace epv
