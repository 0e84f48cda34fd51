// endcond.cuh -- rejection-free exact sampling of a 2-state CTMC path on [0,T] given
// both end states.  Same law as the reference's end_cond_sample_forward_rejection
// (EndCondSampling.cpp:466-509: forward-simulate with holding rates (rate0, rate1)
// and retry until the end state matches), i.e. the CTMC path conditioned on its end
// points (Hobolth & Stone 2009), but drawn jump by jump from the exact conditional
// of the next jump time, so there is no data-dependent retry storm on SIMT:
//   cur == end : stop with prob  exp(-r_cur u) / P_cur,cur(u);  otherwise the next jump
//                has density  ~ exp(-r_cur t) (1 - exp(-s (u-t)))  on (0,u)
//                (truncated-exponential proposal, acceptance >= ~1/2);
//   cur != end : the next jump has density  ~ r_cur exp(-r_cur t) + r_oth exp(-s u) exp(r_oth t),
//                an exact two-component mixture of truncated exponentials.
// u = remaining time, s = rate0 + rate1.
#pragma once
#include "common.cuh"

namespace epv {

// -log(1 - U (1 - e)) / r : exponential(r) truncated to (0, u) with e = exp(-r u)
__device__ __forceinline__ double
trunc_exp(double r, double one_minus_e, double U) {
  return -log1p(-U*one_minus_e)/r;
}

// Appends the sampled jump times (offset by t_off) to out[n..cap).  Returns the new
// count, or -1 if cap was exceeded (the caller restarts with a larger buffer), or -2
// if the iteration cap was hit.
static __device__ __noinline__ int
end_cond_sample(double rate0, double rate1, int start, int end, double T,
                double t_off, Philox &rng, double *out, int n, int cap,
                bool first_jump_given = false) {
  const double s = rate0 + rate1;
  int cur = start;
  double t = 0.0;
  for (int iter = 0; iter < (1 << 20); ++iter) {
    const double u = T - t;
    const double ra = cur ? rate1 : rate0, rb = cur ? rate0 : rate1;
    double tau;
    if (cur == end) {
      if (!(u > 0.0)) return n;
      const double ea = exp(-ra*u), es = exp(-s*u);
      const double p_nojump = ea*s/(rb + ra*es);
      // first_jump_given: the caller already drew "at least one jump" for this interval
      if (!(first_jump_given && iter == 0) && rng.uniform() < p_nojump) return n;
      const double ome_a = -expm1(-ra*u), ome_s = -expm1(-s*u);
      int tries = 0;
      for (;;) {
        tau = trunc_exp(ra, ome_a, rng.uniform());
        const double acc = -expm1(-s*(u - tau))/ome_s;
        if (rng.uniform() < acc) break;
        if (++tries > 4096) return -2;
      }
    }
    else {
      if (!(u > 0.0)) return -2;
      const double ome_a = -expm1(-ra*u), ome_b = -expm1(-rb*u);
      const double w1 = ome_a, w2 = (1.0 - ome_a)*ome_b;
      const double pick = rng.uniform()*(w1 + w2);
      const double U = rng.uniform();
      tau = (pick < w1) ? trunc_exp(ra, ome_a, U) : u - trunc_exp(rb, ome_b, U);
    }
    if (!(tau < u)) tau = u*(1.0 - 1.1102230246251565e-16);
    if (!(tau > 0.0)) tau = u*1.1102230246251565e-16;
    t += tau;
    if (n >= cap) return -1;
    out[n++] = t_off + t;
    cur ^= 1;
  }
  return -2;
}

}  // namespace epv
