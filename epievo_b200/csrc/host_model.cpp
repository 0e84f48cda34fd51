// host_model.cpp -- O(8) host-side model arithmetic behind the C-ABI: parameter file
// values <-> triplet rates, rate normalisation, and the MCEM M-step (gradient ascent
// on log-rates under the 5-free-parameter constraint, optional branch-length update).
// Host code on purpose: the inputs are the 16*(n_nodes) J/D doubles the device
// reduced; nothing here touches paths.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <vector>

#include "../../include/epievo_b200.h"

namespace {

struct M22 { double a00, a01, a10, a11; };

// Gibbs pair potentials from horizontal transition probs (EpiEvoModel.cpp:86-96).
M22 potentials_of(const M22 &T) {
  const double x = std::sqrt(T.a01*T.a10);
  return M22{1.0 - T.a01, x, x, 1.0 - T.a10};
}

// Horizontal transition probs implied by the triplet rates (EpiEvoModel.cpp:102-148):
// potentials up to scale from the death/birth and expand/contract ratios, then the
// stochastic matrix whose stationary pair measure they define.
M22 horizontal_of(const double *r) {
  const double root = std::sqrt(r[2]/r[0]);
  const double q00 = 1.0*root, q11 = 1.0*root*(r[1]/r[3]), q01 = 1.0, q10 = 1.0;
  const double delta = std::sqrt(std::pow(q00 - q11, 2) + 4*q01*q10);
  const double den = q00 + q11 + delta;
  M22 T;
  T.a11 = 2*q11/den;
  T.a00 = 2*q00/den;
  T.a01 = 1.0 - T.a00;
  T.a10 = 1.0 - T.a11;
  return T;
}

void put(const M22 &m, double *o) { o[0] = m.a00; o[1] = m.a01; o[2] = m.a10; o[3] = m.a11; }

double llh8(const double *J, const double *D, const double *r) {
  double x = 0;
  for (int i = 0; i < 8; ++i) x += J[i]*std::log(r[i]) - D[i]*r[i];
  return x;
}

// Gradient w.r.t. log-rates with lambda_100=lambda_001, lambda_110=lambda_011 and
// lambda_111 determined by the others (ParamEstimation.cpp:147-184).
void grad8(const double *J, const double *D, const double *r, double *g) {
  const double f = J[7] - D[7]*r[7];
  g[0] = 0.0 + (J[0] - D[0]*r[0] + f);
  g[2] = 0.0 + (J[2] - D[2]*r[2] - f);
  g[1] = 0.0 + (J[1] + J[4] - (D[1] + D[4])*r[1] - 2*f);
  g[4] = g[1];
  g[3] = 0.0 + (J[3] + J[6] - (D[3] + D[6])*r[3] + 2*f);
  g[6] = g[3];
  g[5] = 0.0 + (J[5] - D[5]*r[5] + f);
  g[7] = 0.0;
}

// Backtracking gradient ascent until no step improves (ParamEstimation.cpp:200-296).
double ascend(double tol, const double *J, const double *D, const double *start,
              double *best) {
  double cur[8], cand[8];
  std::memcpy(cur, start, sizeof cur);
  double llh = llh8(J, D, cur);
  for (;;) {
    double g[8], norm = 0.0;
    grad8(J, D, cur, g);
    for (int i = 0; i < 8; ++i) norm += std::fabs(g[i]);
    double step = 1.0/norm, cand_llh = -DBL_MAX;
    while (cand_llh < llh && step > tol) {
      for (int i = 0; i < 7; ++i) cand[i] = std::exp(std::log(cur[i]) + g[i]*step);
      cand[7] = std::exp(std::log(cand[0]) + std::log(cand[5]) + 2*std::log(cand[3])
                         - std::log(cand[2]) - 2*std::log(cand[1]));
      cand_llh = llh8(J, D, cand);
      step *= 0.5;
    }
    if (!(cand_llh > llh)) break;
    llh = cand_llh;
    std::memcpy(cur, cand, sizeof cur);
  }
  std::memcpy(best, cur, sizeof cur);
  return llh;
}

void collapse(uint64_t n_nodes, const double *X, double *out, const double *w = nullptr) {
  for (int i = 0; i < 8; ++i) out[i] = 0.0;
  for (uint64_t b = 1; b < n_nodes; ++b)
    for (int i = 0; i < 8; ++i) out[i] += w ? w[b]*X[b*8 + i] : X[b*8 + i];
}

}  // namespace

extern "C" {

double epv_rate_scaling_factor(const double *rates) {
  const M22 T = horizontal_of(rates);
  const double t[4] = {T.a00, T.a01, T.a10, T.a11};
  double pi[2];
  pi[1] = (1.0 - T.a00)/(2.0 - T.a00 - T.a11);
  pi[0] = 1.0 - pi[1];
  double mu = 0.0;
  for (int i = 0; i < 8; ++i) {
    const int l = (i >> 2) & 1, m = (i >> 1) & 1, r = i & 1;
    mu += pi[l]*t[l*2 + m]*t[m*2 + r]*rates[i];
  }
  return mu;
}

void epv_model_rebuild(const double *rates, double *T, double *Q, double *baseline) {
  const M22 t = horizontal_of(rates);
  const M22 q = potentials_of(t);
  put(t, T); put(q, Q);
  const double l01 = std::log(q.a01), l10 = std::log(q.a10), l11 = std::log(q.a11);
  baseline[0] = std::log(rates[0]) - (l01 + l10);
  baseline[1] = std::log(rates[1]) - (l01 + l11);
  baseline[2] = std::log(rates[4]) - (l11 + l10);
  baseline[3] = std::log(rates[7]) - (l10 + l01);
  const double centre = baseline[1];
  for (int i = 0; i < 4; ++i) baseline[i] -= centre;
}

void epv_model_from_params(double T00, double T11, double b00, double b11, int scale,
                           double *rates, double *T, double *Q, double *baseline) {
  M22 t; t.a00 = T00; t.a11 = T11; t.a10 = 1.0 - T11; t.a01 = 1.0 - T00;
  const M22 q = potentials_of(t);
  put(t, T); put(q, Q);
  baseline[0] = b00; baseline[1] = 0.0; baseline[2] = 0.0; baseline[3] = b11;
  // rates of pattern l m r: potentials of the two pairs after the flip times the
  // exponentiated baseline of the (l, r) context (EpiEvoModel.cpp:398-417)
  const double e00 = std::exp(b00), e01 = std::exp(0.0), e10 = std::exp(0.0), e11 = std::exp(b11);
  rates[0] = q.a01*q.a10*e00;  rates[2] = q.a00*q.a00*e00;
  rates[1] = q.a01*q.a11*e01;  rates[3] = q.a00*q.a01*e01;
  rates[4] = q.a11*q.a10*e10;  rates[6] = q.a10*q.a00*e10;
  rates[5] = q.a11*q.a11*e11;  rates[7] = q.a10*q.a01*e11;
  if (scale) {
    const double mu = epv_rate_scaling_factor(rates);
    for (int i = 0; i < 8; ++i) rates[i] /= mu;
  }
}

double epv_estimate_rates(uint64_t n_nodes, const double *J, const double *D,
                          double param_tol, double *rates, double *T, double *baseline) {
  double Jc[8], Dc[8], best[8], Q[4];
  collapse(n_nodes, J, Jc); collapse(n_nodes, D, Dc);
  const double llh = ascend(param_tol, Jc, Dc, rates, best);
  std::memcpy(rates, best, sizeof best);
  epv_model_rebuild(rates, T, Q, baseline);
  return llh;
}

void epv_set_one_change(double *rates, uint64_t n_nodes, double *branches) {
  const double mu = epv_rate_scaling_factor(rates);
  for (uint64_t b = 0; b < n_nodes; ++b) branches[b] = branches[b]*mu;
  for (int i = 0; i < 8; ++i) rates[i] = rates[i]/mu;
}

double epv_estimate_rates_and_branches(uint64_t n_nodes, const double *J, const double *D,
                                       double param_tol, double *rates, double *branches,
                                       double *T, double *baseline) {
  double Jc[8], Dc[8], best[8], Q[4];
  collapse(n_nodes, J, Jc); collapse(n_nodes, D, Dc);
  ascend(param_tol, Jc, Dc, rates, best);
  // per-branch time scale = expected / observed jumps (ParamEstimation.cpp:224-240)
  std::vector<double> scale(n_nodes, 1.0), nb(n_nodes);
  for (uint64_t b = 1; b < n_nodes; ++b) {
    double num = 0.0, den = 0.0;
    for (int i = 0; i < 8; ++i) num += J[b*8 + i];
    for (int i = 0; i < 8; ++i) den += D[b*8 + i]*best[i];
    scale[b] = num/den;
  }
  for (uint64_t b = 0; b < n_nodes; ++b) nb[b] = scale[b]*branches[b];
  epv_set_one_change(best, n_nodes, nb.data());
  epv_model_rebuild(best, T, Q, baseline);
  collapse(n_nodes, D, Dc, scale.data());
  std::memcpy(rates, best, sizeof best);
  std::memcpy(branches, nb.data(), n_nodes*sizeof(double));
  return llh8(Jc, Dc, rates);
}

}  // extern "C"

/* ---- site-independent 2-rate model of epievo_initialization (IndepSite.cpp) ------------ */
extern "C" {

/* Per-branch tables of the 2-state CTMC with rates (r0, r1) over time T: transition
 * probabilities (continuous_time_trans_prob_mat, ContinuousTimeMarkovModel.cpp:142-161, which
 * forms h as 1/exp(+x)) and the end-conditioned expectations of jump counts and holding times
 * (expectation_J / expectation_D, :167-226).  2x2 matrices row-major over (start, end). */
void epv_indep_tables(double r0, double r1, double T, double *P, double *J0, double *J1,
                      double *D0, double *D1) {
  const double s = r0 + r1, p = r0*r1;
  {
    const double h = 1.0/std::exp(T*s);
    P[0] = (r0*h + r1)/s; P[1] = 1.0 - P[0];
    P[3] = (r0 + r1*h)/s; P[2] = 1.0 - P[3];
  }
  const double e = std::exp(-s*T);
  {
    const double d = r1 - r0;
    const double C1 = d*(1 - e)/s;
    J0[0] = p*(T*(r1 - r0*e) - C1)/(s*(r1 + r0*e)); J1[0] = J0[0];
    J0[3] = p*(T*(r0 - r1*e) + C1)/(s*(r0 + r1*e)); J1[3] = J0[3];
    const double C2 = p*T*(1 + e)/(s*(1 - e));
    const double C3 = (r0*r0 + r1*r1)/(s*s), C4 = (2*p)/(s*s);
    J0[1] = C2 + C3; J1[1] = C2 - C4;
    J0[2] = J1[1]; J1[2] = J0[1];
  }
  {
    const double r00 = r0*r0, r11 = r1*r1;
    const double C1 = 2*p*(1 - e)/s;
    D0[0] = ((r11 + r00*e)*T + C1)/(s*(r1 + r0*e)); D1[0] = T - D0[0];
    D1[3] = ((r00 + r11*e)*T + C1)/(s*(r0 + r1*e)); D0[3] = T - D1[3];
    const double C2 = (p - r00)*(1 - e)/s;
    D1[1] = ((r00 - p*e)*T + C2)/(s*(r0 - r0*e)); D0[1] = T - D1[1];
    const double C3 = (p - r11)*(1 - e)/s;
    D0[2] = ((r11 - p*e)*T + C3)/(s*(r1 - r1*e)); D1[2] = T - D0[2];
  }
}

/* estimate_rates_indep (IndepSite.cpp:306-326): J, D are [n_nodes][2], entry 0 unused. */
void epv_indep_estimate_rates(uint64_t n_nodes, const double *J, const double *D, double *rates) {
  double Js[2] = {0.0, 0.0}, Ds[2] = {0.0, 0.0};
  for (uint64_t b = 1; b < n_nodes; ++b) {
    Js[0] += J[2*b]; Js[1] += J[2*b + 1];
    Ds[0] += D[2*b]; Ds[1] += D[2*b + 1];
  }
  if (Ds[0] > 0) rates[0] = std::max(Js[0]/Ds[0], 10e-6);
  if (Ds[1] > 0) rates[1] = std::max(Js[1]/Ds[1], 10e-6);
}

/* estimate_rates_and_branches_indep (:341-374) without its last loop: the caller rescales
 * the stored paths with epv_scale_jump_times(ctx, branches). */
void epv_indep_estimate_rates_and_branches(uint64_t n_nodes, const double *J, const double *D,
                                           double *rates, double *branches) {
  epv_indep_estimate_rates(n_nodes, J, D, rates);
  for (uint64_t b = 1; b < n_nodes; ++b)
    branches[b] *= (J[2*b] + J[2*b + 1])/(D[2*b]*rates[0] + D[2*b + 1]*rates[1]);
  const double pi0 = rates[1]/(rates[0] + rates[1]), pi1 = rates[0]/(rates[0] + rates[1]);
  const double scale = pi0*rates[0] + pi1*rates[1];                 /* indep_rate_scaling_factor */
  for (uint64_t b = 0; b < n_nodes; ++b) branches[b] *= scale;
  rates[0] /= scale; rates[1] /= scale;
}

}  // extern "C"
