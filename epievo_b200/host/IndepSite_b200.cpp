/* IndepSite_b200.cpp -- drop-in for the reference's IndepSite.o (src/libepievo/IndepSite.cpp,
 * interface IndepSite.hpp:38-72) over the C-ABI of libepievo_b200.so: the O(n_sites) parts
 * (expectation_sufficient_statistics, update_paths_indep, compute_sufficient_statistics) run
 * on the GPU, the O(n_nodes) M-step on the host through the library's host functions.
 * prog/epievo_initialization.cpp compiles unchanged and links this object instead of
 * IndepSite.o (integration/Makefile).  Host glue only: flatten vector<vector<Path>>
 * [site][node] into the node-major CSR exchange format and back.
 *
 * Randomness: update_paths_indep takes two 32-bit draws from the caller's mt19937 as the
 * Philox seed (the generator still advances, -s still makes runs reproducible); draw-by-draw
 * the histories differ from the CPU build, their law is the same
 * (tests/test_gpu_indep.py). */
#include <random>       /* IndepSite.hpp names std::mt19937 without including it */

#include "IndepSite.hpp"

#include <cstdlib>
#include <stdexcept>
#include <string>

#include "epievo_b200.h"

using std::vector;

namespace {

struct Device {
  epv_ctx *ctx;
  uint64_t iteration;
  Device() : ctx(nullptr), iteration(0) {
    const char *e = std::getenv("EPIEVO_B200_DEVICE");
    if (epv_create(&ctx, e ? std::atoi(e) : 0) != 0 || !ctx)
      throw std::runtime_error("epievo_b200: no usable CUDA device (there is no CPU fallback)");
  }
  ~Device() { epv_destroy(ctx); }
  void check(int rc) const {
    if (rc != 0) throw std::runtime_error(std::string("epievo_b200: ") + epv_last_error(ctx));
  }
};

Device &device() { static Device d; return d; }

struct Flat {
  vector<uint8_t> init;
  vector<uint64_t> off;
  vector<double> jumps, tot;
};

void
upload(Device &d, const vector<vector<Path> > &paths, const vector<uint64_t> &subtree,
       const vector<uint64_t> &parent, Flat &f) {
  const size_t n_sites = paths.size(), n_nodes = paths[0].size();
  f.init.assign(n_sites*n_nodes, 0);
  f.off.assign(n_sites*n_nodes + 1, 0);
  f.tot.assign(n_nodes, 0.0);
  size_t total = 0;
  for (size_t s = 0; s < n_sites; ++s)
    for (size_t b = 1; b < n_nodes; ++b) total += paths[s][b].jumps.size();
  f.jumps.resize(total ? total : 1);
  size_t k = 0;
  for (size_t b = 1; b < n_nodes; ++b) {
    f.tot[b] = paths[0][b].tot_time;
    for (size_t s = 0; s < n_sites; ++s) {
      const Path &p = paths[s][b];
      const size_t r = b*n_sites + s;
      f.init[r] = p.init_state;
      f.off[r] = k;
      for (size_t j = 0; j < p.jumps.size(); ++j) f.jumps[k++] = p.jumps[j];
    }
  }
  f.off[n_sites*n_nodes] = k;
  d.check(epv_set_tree(d.ctx, n_nodes, subtree.data(), parent.data(), f.tot.data()));
  d.check(epv_upload_paths(d.ctx, n_sites, n_nodes, f.init.data(), f.off.data(), f.jumps.data(),
                           f.tot.data()));
}

/* When no TreeHelper is passed (compute_sufficient_statistics) any valid topology will do:
 * the per-row statistics do not look at it.  A caterpillar over n_nodes nodes. */
void
any_tree(size_t n_nodes, vector<uint64_t> &subtree, vector<uint64_t> &parent) {
  subtree.resize(n_nodes); parent.resize(n_nodes);
  for (size_t v = 0; v < n_nodes; ++v) { subtree[v] = v == 0 ? n_nodes : 1; parent[v] = 0; }
}

void
tree_arrays(const TreeHelper &th, vector<uint64_t> &subtree, vector<uint64_t> &parent) {
  subtree.assign(th.subtree_sizes.begin(), th.subtree_sizes.end());
  parent.assign(th.parent_ids.begin(), th.parent_ids.end());
}

void
unflatten2(const vector<double> &X, size_t n_nodes, double div, vector<vector<double> > &out,
           bool row0) {
  out.assign(n_nodes, vector<double>(2, 0.0));
  if (!row0) out[0].clear();                 /* the reference leaves J[0], D[0] empty (:274-281) */
  for (size_t b = 1; b < n_nodes; ++b) { out[b][0] = X[2*b]/div; out[b][1] = X[2*b + 1]/div; }
}

}  // namespace

/* IndepSite.cpp:221-239 */
void
expectation_sufficient_statistics(const vector<double> &rates, const TreeHelper &th,
                                  const vector<vector<Path> > &paths,
                                  vector<vector<double> > &J, vector<vector<double> > &D) {
  Device &d = device();
  vector<uint64_t> ss, pi; tree_arrays(th, ss, pi);
  Flat f; upload(d, paths, ss, pi, f);
  vector<double> Jf(2*th.n_nodes), Df(2*th.n_nodes);
  d.check(epv_indep_expectation(d.ctx, rates.data(), Jf.data(), Df.data()));
  unflatten2(Jf, th.n_nodes, 1.0, J, true);
  unflatten2(Df, th.n_nodes, 1.0, D, true);
}

/* IndepSite.cpp:243-260 */
void
update_paths_indep(const vector<double> &rates, const TreeHelper &th,
                   vector<vector<Path> > &paths, std::mt19937 &gen) {
  Device &d = device();
  vector<uint64_t> ss, pi; tree_arrays(th, ss, pi);
  Flat f; upload(d, paths, ss, pi, f);
  const uint64_t seed_hi = gen();
  const uint64_t seed_lo = gen();
  d.check(epv_indep_update_paths(d.ctx, rates.data(), (seed_hi << 32) | seed_lo, d.iteration++));
  uint64_t n_sites = 0, n_nodes = 0, n_jumps = 0;
  d.check(epv_paths_size(d.ctx, &n_sites, &n_nodes, &n_jumps));
  f.jumps.resize(n_jumps ? n_jumps : 1);
  d.check(epv_download_paths(d.ctx, f.init.data(), f.off.data(), f.jumps.data(), f.tot.data()));
  for (size_t s = 0; s < n_sites; ++s) {
    paths[s][0] = Path(f.init[n_sites + s] != 0, 0.0);       /* sampled_path.front(), :186 */
    for (size_t b = 1; b < n_nodes; ++b) {
      const size_t r = b*n_sites + s;
      Path &p = paths[s][b];
      p.init_state = f.init[r] != 0;
      p.tot_time = f.tot[b];
      p.jumps.assign(f.jumps.begin() + f.off[r], f.jumps.begin() + f.off[r + 1]);
    }
  }
}

/* IndepSite.cpp:266-303 */
void
compute_sufficient_statistics(const vector<vector<Path> > &paths,
                              vector<vector<double> > &J, vector<vector<double> > &D) {
  Device &d = device();
  const size_t n_nodes = paths[0].size();
  vector<uint64_t> ss, pi; any_tree(n_nodes, ss, pi);
  Flat f; upload(d, paths, ss, pi, f);
  vector<double> Jf(2*n_nodes), Df(2*n_nodes);
  d.check(epv_indep_suffstats(d.ctx, Jf.data(), Df.data()));
  unflatten2(Jf, n_nodes, (double)paths.size(), J, false);
  unflatten2(Df, n_nodes, (double)paths.size(), D, false);
}

/* IndepSite.cpp:306-326 */
void
estimate_rates_indep(const vector<vector<double> > &J, const vector<vector<double> > &D,
                     vector<double> &rates, TreeHelper &th) {
  vector<double> Jf(2*th.n_nodes, 0.0), Df(2*th.n_nodes, 0.0);
  for (size_t b = 1; b < th.n_nodes; ++b) {
    Jf[2*b] = J[b][0]; Jf[2*b + 1] = J[b][1]; Df[2*b] = D[b][0]; Df[2*b + 1] = D[b][1];
  }
  epv_indep_estimate_rates(th.n_nodes, Jf.data(), Df.data(), rates.data());
}

/* IndepSite.cpp:341-374 */
void
estimate_rates_and_branches_indep(const vector<vector<double> > &J,
                                  const vector<vector<double> > &D, vector<double> &rates,
                                  TreeHelper &th, vector<vector<Path> > &paths) {
  vector<double> Jf(2*th.n_nodes, 0.0), Df(2*th.n_nodes, 0.0);
  for (size_t b = 1; b < th.n_nodes; ++b) {
    Jf[2*b] = J[b][0]; Jf[2*b + 1] = J[b][1]; Df[2*b] = D[b][0]; Df[2*b + 1] = D[b][1];
  }
  epv_indep_estimate_rates_and_branches(th.n_nodes, Jf.data(), Df.data(), rates.data(),
                                        th.branches.data());
  for (size_t b = 1; b < paths[0].size(); ++b)             /* scale jump times, :364-372 */
    for (size_t i = 0; i < paths.size(); ++i) {
      const double scale = th.branches[b]/paths[i][b].tot_time;
      for (size_t j = 0; j < paths[i][b].jumps.size(); ++j) paths[i][b].jumps[j] *= scale;
      paths[i][b].tot_time = th.branches[b];
    }
}

/* IndepSite.cpp:377-392 */
void
estimate_root_stationary(const vector<vector<Path> > &paths, vector<double> &pi) {
  pi.resize(2);
  double N1 = 0.0;
  for (size_t s = 0; s < paths.size(); ++s) N1 += paths[s][1].init_state;
  pi[1] = N1/paths.size();
  pi[0] = 1 - pi[1];
}
