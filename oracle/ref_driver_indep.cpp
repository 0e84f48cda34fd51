/* ref_driver_indep.cpp -- TEST INFRASTRUCTURE ONLY.
 * extern "C" driver around the UNMODIFIED reference's site-independent E-step
 * (/root/reference/src/libepievo/IndepSite.cpp), compiled by oracle/Makefile into
 * oracle/_ref/libepievo_ref.so next to ref_driver.cpp.  A separate translation unit because
 * IndepSite.hpp and SingleSiteSampler.cpp (textually included by ref_driver.cpp) both define
 * a global `struct FelsHelper`. */
#include <cstdint>
#include <random>
#include <vector>

#include "IndepSite.hpp"

using std::vector;

namespace {

vector<vector<Path> > i_paths;   // [site][node]

TreeHelper
tree_of(const size_t n_nodes, const uint64_t *subtree_sizes, const uint64_t *parent_ids,
        const double *branches) {
  TreeHelper th;
  th.n_nodes = n_nodes;
  th.subtree_sizes.assign(subtree_sizes, subtree_sizes + n_nodes);
  th.parent_ids.assign(parent_ids, parent_ids + n_nodes);
  th.branches.assign(branches, branches + n_nodes);
  th.node_names.resize(n_nodes);
  return th;
}

void
flatten(const vector<vector<double> > &X, const size_t n_nodes, double *out) {
  for (size_t b = 0; b < n_nodes; ++b)
    for (size_t k = 0; k < 2; ++k)
      out[2*b + k] = (b < X.size() && k < X[b].size()) ? X[b][k] : 0.0;
}

}  // namespace

extern "C" {

void
ref_indep_load(uint64_t n_sites, uint64_t n_nodes, const uint8_t *init_state,
               const uint64_t *row_off, const double *jumps, const double *tot_time) {
  i_paths.assign(n_sites, vector<Path>(n_nodes));
  for (size_t b = 1; b < n_nodes; ++b)
    for (size_t s = 0; s < n_sites; ++s) {
      const size_t row = b*n_sites + s;
      Path &p = i_paths[s][b];
      p.init_state = init_state[row] != 0;
      p.tot_time = tot_time[b];
      p.jumps.assign(jumps + row_off[row], jumps + row_off[row + 1]);
    }
}

uint64_t
ref_indep_total_jumps() {
  uint64_t n = 0;
  for (size_t s = 0; s < i_paths.size(); ++s)
    for (size_t b = 1; b < i_paths[s].size(); ++b) n += i_paths[s][b].jumps.size();
  return n;
}

void
ref_indep_store(uint8_t *init_state, uint64_t *row_off, double *jumps, double *tot_time) {
  const size_t n_sites = i_paths.size(), n_nodes = n_sites ? i_paths[0].size() : 0;
  uint64_t off = 0;
  for (size_t b = 0; b < n_nodes; ++b) {
    tot_time[b] = (b && n_sites) ? i_paths[0][b].tot_time : 0.0;
    for (size_t s = 0; s < n_sites; ++s) {
      const size_t row = b*n_sites + s;
      const Path &p = i_paths[s][b];
      init_state[row] = p.init_state;
      row_off[row] = off;
      for (size_t k = 0; k < p.jumps.size(); ++k) jumps[off++] = p.jumps[k];
    }
  }
  row_off[n_nodes*n_sites] = off;
}

/* compute_sufficient_statistics, IndepSite.cpp:266-303 (averages over sites) */
void
ref_indep_suffstats(uint64_t n_nodes, double *J, double *D) {
  vector<vector<double> > Jv, Dv;
  compute_sufficient_statistics(i_paths, Jv, Dv);
  flatten(Jv, n_nodes, J); flatten(Dv, n_nodes, D);
}

/* expectation_sufficient_statistics, IndepSite.cpp:221-239 (sums over sites) */
void
ref_indep_expectation(uint64_t n_nodes, const uint64_t *subtree_sizes, const uint64_t *parent_ids,
                      const double *branches, const double *rates, double *J, double *D) {
  const TreeHelper th = tree_of(n_nodes, subtree_sizes, parent_ids, branches);
  const vector<double> r(rates, rates + 2);
  vector<vector<double> > Jv, Dv;
  expectation_sufficient_statistics(r, th, i_paths, Jv, Dv);
  flatten(Jv, n_nodes, J); flatten(Dv, n_nodes, D);
}

/* update_paths_indep, IndepSite.cpp:243-260, with a freshly seeded mt19937 */
void
ref_indep_update_paths(uint64_t n_nodes, const uint64_t *subtree_sizes, const uint64_t *parent_ids,
                       const double *branches, const double *rates, uint64_t seed) {
  const TreeHelper th = tree_of(n_nodes, subtree_sizes, parent_ids, branches);
  const vector<double> r(rates, rates + 2);
  std::mt19937 gen(seed);
  update_paths_indep(r, th, i_paths, gen);
}

/* estimate_rates_indep (:306-326) and estimate_rates_and_branches_indep (:341-374, which also
 * rescales the loaded paths) */
void
ref_indep_estimate(uint64_t n_nodes, const uint64_t *subtree_sizes, const uint64_t *parent_ids,
                   double *branches, const double *J, const double *D, double *rates,
                   int with_branches) {
  TreeHelper th = tree_of(n_nodes, subtree_sizes, parent_ids, branches);
  vector<vector<double> > Jv(n_nodes, vector<double>(2)), Dv(n_nodes, vector<double>(2));
  for (size_t b = 0; b < n_nodes; ++b)
    for (size_t k = 0; k < 2; ++k) { Jv[b][k] = J[2*b + k]; Dv[b][k] = D[2*b + k]; }
  vector<double> r(rates, rates + 2);
  if (with_branches) estimate_rates_and_branches_indep(Jv, Dv, r, th, i_paths);
  else estimate_rates_indep(Jv, Dv, r, th);
  rates[0] = r[0]; rates[1] = r[1];
  for (size_t b = 0; b < n_nodes; ++b) branches[b] = th.branches[b];
}

}  // extern "C"
