/* oracle/ref_driver.cpp -- TEST INFRASTRUCTURE ONLY.
 *
 * A thin extern "C" driver around the UNMODIFIED reference sources under
 * /root/reference/src/libepievo.  It is compiled by oracle/Makefile from the
 * sources where they lie (nothing is copied into this repository) into
 * oracle/_ref/libepievo_ref.so, and is used by tests/ (and by bench.py's
 * cpu_baseline / --impl reference legs) as the checker and CPU baseline.
 * The product path (epievo_b200/) never links, loads or calls this.
 *
 * The translation unit includes SingleSiteSampler.cpp textually so that the
 * file-local helpers (FelsHelper, pruning_branch, proposal_prob_branch, ...)
 * are reachable; everything else is linked from the reference objects.
 *
 * Flat "node-major CSR" exchange format used by every entry point
 * (the same one the product C-ABI uses, include/epievo_b200.h):
 *   init_state : uint8  [n_nodes * n_sites]      row = node*n_sites + site
 *   row_off    : uint64 [n_nodes * n_sites + 1]  offsets into jumps
 *   jumps      : double [row_off[last]]
 *   tot_time   : double [n_nodes]                branch length carried by the rows
 * Node 0 (the root) has empty rows, as in the reference's transposed
 * paths[site][node] container (epievo_est_params_histories.cpp:180-192).
 */
#include <cstdint>
#include <cstring>
#include <chrono>

#include "SingleSiteSampler.cpp"   // reference TU, included for its statics
#include "ContinuousTimeMarkovModel.hpp"

namespace {

typedef vector<vector<Path> > PathMat;   // [site][node]

PathMat g_paths;
std::mt19937 g_gen(0);

void
flat_to_paths(const size_t n_sites, const size_t n_nodes,
              const uint8_t *init_state, const uint64_t *row_off,
              const double *jumps, const double *tot_time, PathMat &paths) {
  paths.assign(n_sites, vector<Path>(n_nodes));
  for (size_t b = 1; b < n_nodes; ++b)
    for (size_t s = 0; s < n_sites; ++s) {
      const size_t row = b*n_sites + s;
      Path &p = paths[s][b];
      p.init_state = init_state[row] != 0;
      p.tot_time = tot_time[b];
      p.jumps.assign(jumps + row_off[row], jumps + row_off[row + 1]);
    }
}

TreeHelper
make_tree(const size_t n_nodes, const uint64_t *subtree_sizes,
          const uint64_t *parent_ids, const double *branches) {
  TreeHelper th;
  th.n_nodes = n_nodes;
  th.subtree_sizes.assign(subtree_sizes, subtree_sizes + n_nodes);
  th.parent_ids.assign(parent_ids, parent_ids + n_nodes);
  th.branches.assign(branches, branches + n_nodes);
  th.node_names.resize(n_nodes);
  return th;
}

EpiEvoModel
make_model(const double *rates, const double *T) {
  EpiEvoModel m;
  for (size_t i = 0; i < 8; ++i) m.triplet_rates[i] = rates[i];
  m.T = two_by_two(T[0], T[1], T[2], T[3]);
  return m;
}

void
copy22(const two_by_two &m, double *out) {
  out[0] = m(0, 0); out[1] = m(0, 1); out[2] = m(1, 0); out[3] = m(1, 1);
}

Path
make_path(const int init, const double tot, const size_t n, const double *j) {
  return Path(init != 0, tot, vector<double>(j, j + n));
}

}  // namespace

extern "C" {

/* ---- a3: 2-state CTMC closed forms -------------------------------------- */
void
ref_trans_prob_mat(double r0, double r1, double t, double *P) {
  two_by_two M;
  continuous_time_trans_prob_mat(r0, r1, t, M);
  copy22(M, P);
}

double
ref_trans_prob(double r0, double r1, double t, int a, int b) {
  return TwoStateCTMarkovModel(r0, r1).get_trans_prob(t, a, b);
}

void
ref_expectations(double r0, double r1, double T,
                 double *J0, double *J1, double *D0, double *D1) {
  two_by_two a, b, c, d;
  expectation_J(r0, r1, T, a, b);
  expectation_D(r0, r1, T, c, d);
  copy22(a, J0); copy22(b, J1); copy22(c, D0); copy22(d, D1);
}

/* ---- a2: neighbour interval merge ---------------------------------------- */
uint64_t
ref_segment_info(const double *rates, int l_init, uint64_t nl, const double *lj,
                 int r_init, uint64_t nr, const double *rj, double tot_time,
                 double *rate0, double *rate1, uint64_t *trip0, uint64_t *trip1,
                 double *len) {
  array<double, 8> r;
  for (size_t i = 0; i < 8; ++i) r[i] = rates[i];
  vector<SegmentInfo> si;
  collect_segment_info(r, make_path(l_init, tot_time, nl, lj),
                       make_path(r_init, tot_time, nr, rj), si);
  for (size_t i = 0; i < si.size(); ++i) {
    rate0[i] = si[i].rate0; rate1[i] = si[i].rate1;
    trip0[i] = si[i].trip0; trip1[i] = si[i].trip1; len[i] = si[i].len;
  }
  return si.size();
}

/* ---- a12: three-way J/D merge -------------------------------------------- */
void
ref_add_suff_stats(int l_init, uint64_t nl, const double *lj,
                   int m_init, uint64_t nm, const double *mj,
                   int r_init, uint64_t nr, const double *rj,
                   double tot_time, double *J, double *D) {
  vector<double> Jv(J, J + 8), Dv(D, D + 8);
  add_sufficient_statistics(make_path(l_init, tot_time, nl, lj),
                            make_path(m_init, tot_time, nm, mj),
                            make_path(r_init, tot_time, nr, rj), Jv, Dv);
  std::copy(Jv.begin(), Jv.end(), J);
  std::copy(Dv.begin(), Dv.end(), D);
}

/* ---- path container ------------------------------------------------------- */
void
ref_load_paths(uint64_t n_sites, uint64_t n_nodes, const uint8_t *init_state,
               const uint64_t *row_off, const double *jumps,
               const double *tot_time) {
  flat_to_paths(n_sites, n_nodes, init_state, row_off, jumps, tot_time, g_paths);
}

uint64_t
ref_total_jumps() {
  uint64_t n = 0;
  for (size_t s = 0; s < g_paths.size(); ++s)
    for (size_t b = 1; b < g_paths[s].size(); ++b)
      n += g_paths[s][b].jumps.size();
  return n;
}

void
ref_store_paths(uint8_t *init_state, uint64_t *row_off, double *jumps,
                double *tot_time) {
  const size_t n_sites = g_paths.size();
  const size_t n_nodes = n_sites ? g_paths[0].size() : 0;
  uint64_t off = 0;
  for (size_t b = 0; b < n_nodes; ++b) {
    tot_time[b] = (b && n_sites) ? g_paths[0][b].tot_time : 0.0;
    for (size_t s = 0; s < n_sites; ++s) {
      const size_t row = b*n_sites + s;
      const Path &p = g_paths[s][b];
      init_state[row] = p.init_state;
      row_off[row] = off;
      for (size_t k = 0; k < p.jumps.size(); ++k) jumps[off++] = p.jumps[k];
    }
  }
  row_off[n_nodes*n_sites] = off;
}

/* ---- local-paths files: the reference's reader / writer ------- */
/* read_paths (Path.cpp:123-148) + the transposition every caller does
 * (prog/epievo_est_params_histories.cpp:183-192).  Returns seconds spent in read_paths,
 * < 0 on failure. */
double
ref_read_paths_file(const char *filename, uint64_t *n_sites, uint64_t *n_nodes) {
  vector<string> names;
  vector<vector<Path> > orig;
  const auto t0 = std::chrono::steady_clock::now();
  try { read_paths(filename, names, orig); }
  catch (const std::exception &) { return -1.0; }
  const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  if (orig.size() < 2) return -1.0;
  const size_t ns = orig[1].size(), nn = orig.size();
  g_paths.assign(ns, vector<Path>(nn));
  for (size_t i = 0; i < ns; ++i)
    for (size_t b = 1; b < nn; ++b) g_paths[i][b] = orig[b][i];
  *n_sites = ns; *n_nodes = nn;
  return sec;
}

/* The writer of prog/epievo_est_params_histories.cpp:56-75 (static there): the same
 * statements over operator<<(ostream, Path) (Path.cpp:62-71).  Returns seconds. */
double
ref_write_paths_file(const char *filename, uint64_t n_nodes, const char *const *names) {
  const auto t0 = std::chrono::steady_clock::now();
  {
    std::ofstream out(filename);
    if (!out) return -1.0;
    out << "NODE:" << names[0] << std::endl;
  }
  for (size_t b = 1; b < n_nodes; ++b) {
    std::ofstream out(filename, std::ofstream::app);
    if (!out) return -1.0;
    out << "NODE:" << names[b] << std::endl;
    for (size_t i = 0; i < g_paths.size(); ++i)
      out << i << '\t' << g_paths[i][b] << '\n';
  }
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

/* ---- a13: whole-data per-branch J/D --------------------------------------- */
void
ref_get_suff_stats(double *J, double *D) {
  vector<vector<double> > Jv, Dv;
  get_sufficient_statistics(g_paths, Jv, Dv);
  for (size_t b = 0; b < Jv.size(); ++b)
    for (size_t t = 0; t < 8; ++t) {
      J[b*8 + t] = Jv[b][t];
      D[b*8 + t] = Dv[b][t];
    }
}

/* ---- a8/a10: tri_llh of every interior site ------------------------------- */
void
ref_tri_llh(const double *rates, const double *T, double *tri_llh) {
  const EpiEvoModel m = make_model(rates, T);
  array<double, 8> log_rates;
  for (size_t i = 0; i < 8; ++i) log_rates[i] = log(rates[i]);
  const size_t n = g_paths.size();
  for (size_t s = 0; s < n; ++s) tri_llh[s] = 0.0;
  for (size_t s = 1; s + 1 < n; ++s)
    tri_llh[s] = path_log_likelihood(m, g_paths[s-1], g_paths[s], g_paths[s+1],
                                     log_rates);
}

/* ---- a2+a4+a7: deterministic per-site intermediates ------------------------
 * seg_off[n_nodes+1] gives each node's slice of the interval arrays;
 * p is [n_intervals][2], q is [n_nodes][2].                                  */
uint64_t
ref_site_debug(uint64_t n_nodes, const uint64_t *subtree_sizes,
               const uint64_t *parent_ids, const double *branches,
               const double *rates, const double *T, uint64_t site,
               uint64_t cap, uint64_t *seg_off, double *rate0, double *rate1,
               uint64_t *trip0, uint64_t *trip1, double *len,
               double *p, double *q, double *root_p0, double *proposal_logp) {
  const TreeHelper th = make_tree(n_nodes, subtree_sizes, parent_ids, branches);
  const EpiEvoModel m = make_model(rates, T);
  vector<vector<SegmentInfo> > seg_info(n_nodes);
  for (size_t b = 1; b < n_nodes; ++b)
    collect_segment_info(m.triplet_rates, g_paths[site-1][b],
                         g_paths[site+1][b], seg_info[b]);
  vector<FelsHelper> fh;
  pruning(th, site, g_paths, seg_info, fh);

  uint64_t k = 0;
  seg_off[0] = 0;
  for (size_t b = 0; b < n_nodes; ++b) {
    for (size_t i = 0; i < seg_info[b].size(); ++i, ++k) {
      if (k >= cap) return UINT64_MAX;
      rate0[k] = seg_info[b][i].rate0; rate1[k] = seg_info[b][i].rate1;
      trip0[k] = seg_info[b][i].trip0; trip1[k] = seg_info[b][i].trip1;
      len[k] = seg_info[b][i].len;
      p[2*k] = fh[b].p[i][0]; p[2*k + 1] = fh[b].p[i][1];
    }
    seg_off[b + 1] = k;
    q[2*b] = fh[b].q[0]; q[2*b + 1] = fh[b].q[1];
  }
  const bool ls = g_paths[site-1][1].init_state;
  const bool rs = g_paths[site+1][1].init_state;
  *root_p0 = root_post_prob0(ls, rs, m.T, fh[0].q);
  *proposal_logp = proposal_prob(m.triplet_rates, th, ls, rs, m.T, fh, seg_info,
                                 g_paths[site], false);
  return k;
}

/* ---- a11: the sampler itself ---------------------------------------------- */
void
ref_seed(uint64_t seed) { g_gen.seed(seed); }

double
ref_run_mcmc(uint64_t n_nodes, const uint64_t *subtree_sizes,
             const uint64_t *parent_ids, const double *branches,
             const double *rates, const double *T,
             uint64_t burn_in, uint64_t batch, int do_reset,
             double *J, double *D, double *acc_rate, double *seconds) {
  static SingleSiteSampler *mcmc = 0;
  const TreeHelper th = make_tree(n_nodes, subtree_sizes, parent_ids, branches);
  const EpiEvoModel m = make_model(rates, T);
  if (!mcmc || do_reset) {
    delete mcmc;
    mcmc = new SingleSiteSampler(burn_in, batch);
    mcmc->reset(m, g_paths);
  }
  mcmc->burn_in = burn_in;
  mcmc->batch = batch;
  vector<vector<double> > Jv, Dv;
  double acc = 0.0;
  const auto t0 = std::chrono::steady_clock::now();
  mcmc->run_mcmc(m, th, g_paths, g_gen, Jv, Dv, acc);
  const auto t1 = std::chrono::steady_clock::now();
  for (size_t b = 1; b < n_nodes; ++b)
    for (size_t t = 0; t < 8; ++t) {
      J[b*8 + t] = Jv[b][t];
      D[b*8 + t] = Dv[b][t];
    }
  for (size_t t = 0; t < 8; ++t) J[t] = D[t] = 0.0;
  *acc_rate = acc;
  const double sec = std::chrono::duration<double>(t1 - t0).count();
  if (seconds) *seconds = sec;
  return sec;
}

/* Timed J/D pass only (bench cpu_baseline for the suff-stat config). */
double
ref_time_suff_stats(uint64_t reps, double *J, double *D) {
  const auto t0 = std::chrono::steady_clock::now();
  for (uint64_t r = 0; r < reps; ++r) ref_get_suff_stats(J, D);
  const auto t1 = std::chrono::steady_clock::now();
  return std::chrono::duration<double>(t1 - t0).count();
}

/* ---- model helpers (host side of the boundary) ---------------------------- */
void
ref_model_from_params(double T00, double T11, double b00, double b11, int scale,
                      double *rates, double *T, double *Q, double *baseline) {
  EpiEvoModel m;
  m.T = two_by_two(T00, 1.0 - T00, 1.0 - T11, T11);
  m.stationary_baseline = two_by_two(b00, 0.0, 0.0, b11);
  m.initialize();
  if (scale) m.scale_triplet_rates();
  for (size_t i = 0; i < 8; ++i) rates[i] = m.triplet_rates[i];
  copy22(m.T, T); copy22(m.Q, Q); copy22(m.stationary_baseline, baseline);
}

void
ref_model_rebuild(const double *rates_in, double *T, double *Q, double *baseline) {
  EpiEvoModel m;
  array<double, 8> r;
  for (size_t i = 0; i < 8; ++i) r[i] = rates_in[i];
  m.rebuild_from_triplet_rates(r);
  copy22(m.T, T); copy22(m.Q, Q); copy22(m.stationary_baseline, baseline);
}

double
ref_rate_scaling_factor(const double *rates) {
  array<double, 8> r;
  for (size_t i = 0; i < 8; ++i) r[i] = rates[i];
  return rate_scaling_factor(r);
}

/* ---- a15/a16: M-step -------------------------------------------------------- */
static void
unflatten(uint64_t n_nodes, const double *X, vector<vector<double> > &v) {
  v.assign(n_nodes, vector<double>(8, 0.0));
  for (size_t b = 0; b < n_nodes; ++b)
    for (size_t t = 0; t < 8; ++t) v[b][t] = X[b*8 + t];
}

double
ref_estimate_rates(uint64_t n_nodes, const double *J, const double *D,
                   double param_tol, double *rates, double *T, double *baseline) {
  vector<vector<double> > Jv, Dv;
  unflatten(n_nodes, J, Jv); unflatten(n_nodes, D, Dv);
  EpiEvoModel m;
  for (size_t i = 0; i < 8; ++i) m.triplet_rates[i] = rates[i];
  const double llh = estimate_rates(false, param_tol, Jv, Dv, m);
  for (size_t i = 0; i < 8; ++i) rates[i] = m.triplet_rates[i];
  copy22(m.T, T); copy22(m.stationary_baseline, baseline);
  return llh;
}

double
ref_estimate_rates_and_branches(uint64_t n_nodes, const double *J, const double *D,
                                double param_tol, double *rates, double *branches,
                                double *T, double *baseline) {
  vector<vector<double> > Jv, Dv;
  unflatten(n_nodes, J, Jv); unflatten(n_nodes, D, Dv);
  EpiEvoModel m;
  for (size_t i = 0; i < 8; ++i) m.triplet_rates[i] = rates[i];
  TreeHelper th;
  th.n_nodes = n_nodes;
  th.branches.assign(branches, branches + n_nodes);
  const double llh = estimate_rates_and_branches(false, param_tol, Jv, Dv, th, m);
  for (size_t i = 0; i < 8; ++i) rates[i] = m.triplet_rates[i];
  for (size_t b = 0; b < n_nodes; ++b) branches[b] = th.branches[b];
  copy22(m.T, T); copy22(m.stationary_baseline, baseline);
  return llh;
}

void
ref_set_one_change(double *rates, uint64_t n_nodes, double *branches) {
  array<double, 8> r;
  for (size_t i = 0; i < 8; ++i) r[i] = rates[i];
  vector<double> br(branches, branches + n_nodes);
  set_one_change_per_site_per_unit_time(r, br);
  for (size_t i = 0; i < 8; ++i) rates[i] = r[i];
  for (size_t b = 0; b < n_nodes; ++b) branches[b] = br[b];
}

/* ---- a17 --------------------------------------------------------------------- */
void
ref_scale_jump_times(uint64_t n_nodes, const double *new_branches) {
  TreeHelper th;
  th.n_nodes = n_nodes;
  th.branches.assign(new_branches, new_branches + n_nodes);
  scale_jump_times(g_paths, th);
}

/* ---- a6: the hot path's end-conditioned sampler, for law checks -------------- */
uint64_t
ref_end_cond_sample(double r0, double r1, int a, int b, double T,
                    uint64_t cap, double *jumps_out) {
  exponential_distribution<double> e0(r0), e1(r1);
  vector<double> j;
  end_cond_sample_forward_rejection(e0, e1, a, b, T, g_gen, j, 0.0);
  const size_t n = std::min<size_t>(cap, j.size());
  std::copy(j.begin(), j.begin() + n, jumps_out);
  return j.size();
}

}  // extern "C"
