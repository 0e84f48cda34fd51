/* SingleSiteSampler_b200.cpp -- the reference's SingleSiteSampler interface
 * (src/libepievo/SingleSiteSampler.hpp:35-81) implemented over the C-ABI of
 * libepievo_b200.so.  Host glue only: flatten vector<vector<Path>> [site][node] into the
 * node-major CSR exchange format, call the device path, unflatten. */
#include "SingleSiteSampler.hpp"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>
#include <thread>

#include "epievo_b200.h"

using std::vector;

static int
int_from_env(const char *name, int dflt) {
  const char *e = std::getenv(name);
  return e && *e ? std::atoi(e) : dflt;
}

/* EPIEVO_B200_TIMING=1: seconds spent per phase of the drop-in calls, printed to stderr when the sampler
 * is destroyed (tools/cli_bench.py reads it). */
namespace {
struct PhaseClock {
  double create = 0, flatten = 0, upload = 0, device = 0, download = 0, unflatten = 0;
  size_t calls = 0;
  const bool on = std::getenv("EPIEVO_B200_TIMING") != nullptr;
  ~PhaseClock() {
    if (on && calls)
      std::fprintf(stderr, "[epievo_b200 timing] calls=%zu create=%.4f flatten=%.4f upload=%.4f device=%.4f download=%.4f unflatten=%.4f\n",
                   calls, create, flatten, upload, device, download, unflatten);
  }
} phase_clock;
double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
}  // namespace

static epv_multi *
new_context() {
  const int first = int_from_env("EPIEVO_B200_DEVICE", 0), n = std::max(1, int_from_env("EPIEVO_B200_GPUS", 1));
  std::vector<int> dev(n);
  for (int r = 0; r < n; ++r) dev[r] = first + r;
  epv_multi *c = nullptr;
  const double t0 = now_s();                 // (CUDA initialisation + context creation: the process's first CUDA call)
  if (epv_multi_create(&c, n, dev.data()) != 0 || !c)
    throw std::runtime_error("epievo_b200: no usable CUDA device(s) (EPIEVO_B200_DEVICE / EPIEVO_B200_GPUS; "
                             "there is no CPU fallback)");
  phase_clock.create += now_s() - t0;
  return c;
}

SingleSiteSampler::SingleSiteSampler(const size_t n_burn_in, const size_t n_batch) :
  SAMPLE_ROOT(false), burn_in(n_burn_in), batch(n_batch), ctx(new_context()),
  sweep_counter(0), last_sweep_accepts(0) {}

SingleSiteSampler::~SingleSiteSampler() { epv_multi_destroy(ctx); }

void
SingleSiteSampler::check(int rc) const {
  if (rc != 0) throw std::runtime_error(std::string("epievo_b200: ") + epv_multi_last_error(ctx));
}

void
SingleSiteSampler::configure(const EpiEvoModel &the_model, const TreeHelper &th) {
  if (SAMPLE_ROOT)
    throw std::runtime_error("epievo_b200: SAMPLE_ROOT = true is not supported (the device sampler keeps "
                             "the root state, as every caller of the reference does)");
  vector<uint64_t> ss(th.subtree_sizes.begin(), th.subtree_sizes.end());
  vector<uint64_t> pi(th.parent_ids.begin(), th.parent_ids.end());
  check(epv_multi_set_tree(ctx, th.n_nodes, ss.data(), pi.data(), th.branches.data()));
  const double T[4] = {the_model.T(0, 0), the_model.T(0, 1), the_model.T(1, 0), the_model.T(1, 1)};
  check(epv_multi_set_model(ctx, the_model.triplet_rates.data(), T));
}

/* Site ranges handled by the host threads of flatten / unflatten (one heap vector per row makes
 * both memory-latency bound; at 1e6 sites they were the largest term of a run_mcmc call). */
static size_t
host_threads(size_t n_rows) {
  const size_t hw = std::max(1u, std::thread::hardware_concurrency());
  return n_rows < 200000 ? 1 : std::min<size_t>(hw, 32);
}

template <typename F>
static void
for_site_ranges(size_t n_sites, size_t n_threads, F f) {
  if (n_threads <= 1) { f(0, 0, n_sites); return; }
  std::vector<std::thread> th;
  for (size_t t = 0; t < n_threads; ++t)
    th.emplace_back(f, t, n_sites*t/n_threads, n_sites*(t + 1)/n_threads);
  for (size_t t = 0; t < th.size(); ++t) th[t].join();
}

/* [site][node] -> node-major CSR (the order of the local-paths file, Path.cpp:123-148), threaded over
 * site ranges: every thread first counts the jumps of its rows per node, then writes them where they go. */
template <typename Off>
static void
flatten_paths(const vector<vector<Path> > &paths, const size_t n_nodes, vector<unsigned char> &init_state,
              vector<Off> &row_off, vector<double> &jumps, vector<double> &tot_time) {
  const size_t n_sites = paths.size();
  init_state.assign(n_sites*n_nodes, 0);
  row_off.assign(n_sites*n_nodes + 1, 0);
  tot_time.assign(n_nodes, 0.0);
  for (size_t b = 1; b < n_nodes; ++b) tot_time[b] = paths[0][b].tot_time;
  const size_t T = host_threads(n_sites*n_nodes);
  // jumps per (node, thread) -> where every thread's rows of every node start
  vector<size_t> cnt(n_nodes*T, 0);
  for_site_ranges(n_sites, T, [&](size_t t, size_t s0, size_t s1) {
    for (size_t s = s0; s < s1; ++s)
      for (size_t b = 1; b < n_nodes; ++b) cnt[b*T + t] += paths[s][b].jumps.size();
  });
  size_t total = 0;
  for (size_t i = T; i < n_nodes*T; ++i) { const size_t c = cnt[i]; cnt[i] = total; total += c; }
  jumps.resize(total ? total : 1);
  for_site_ranges(n_sites, T, [&](size_t t, size_t s0, size_t s1) {
    for (size_t b = 1; b < n_nodes; ++b) {
      size_t k = cnt[b*T + t];
      for (size_t s = s0; s < s1; ++s) {
        const Path &p = paths[s][b];
        const size_t r = b*n_sites + s;
        init_state[r] = p.init_state;
        row_off[r] = k;
        for (size_t j = 0; j < p.jumps.size(); ++j) jumps[k++] = p.jumps[j];
      }
    }
  });
  row_off[n_sites*n_nodes] = total;
}

void
SingleSiteSampler::upload(const vector<vector<Path> > &paths, const TreeHelper &th) {
  const size_t n_sites = paths.size(), n_nodes = th.n_nodes;
  const double t0 = now_s();
  flatten_paths(paths, n_nodes, init_state, row_off, jumps, tot_time);
  const double t1 = now_s();
  check(epv_multi_upload_paths(ctx, n_sites, n_nodes, init_state.data(),
                               reinterpret_cast<const uint64_t *>(row_off.data()), jumps.data(),
                               tot_time.data()));
  phase_clock.flatten += t1 - t0; phase_clock.upload += now_s() - t1; ++phase_clock.calls;
}

void
SingleSiteSampler::download(vector<vector<Path> > &paths) {
  uint64_t n_sites = 0, n_nodes = 0, n_jumps = 0;
  const double t0 = now_s();
  check(epv_multi_paths_size(ctx, &n_sites, &n_nodes, &n_jumps));
  jumps.resize(n_jumps ? n_jumps : 1);
  check(epv_multi_download_paths(ctx, init_state.data(), reinterpret_cast<uint64_t *>(row_off.data()),
                                 jumps.data(), tot_time.data()));
  const double t1 = now_s();
  phase_clock.download += t1 - t0;
  for_site_ranges(n_sites, host_threads(n_sites*n_nodes), [&](size_t, size_t s0, size_t s1) {
    for (size_t s = s0; s < s1; ++s)
      for (size_t b = 1; b < n_nodes; ++b) {
        const size_t r = b*n_sites + s;
        Path &p = paths[s][b];
        p.init_state = init_state[r] != 0;
        p.tot_time = tot_time[b];
        p.jumps.assign(jumps.begin() + row_off[r], jumps.begin() + row_off[r + 1]);
      }
  });
  phase_clock.unflatten += now_s() - t1;
}

/* The reference caches tri_llh here (SingleSiteSampler.cpp:449-475).  The device cache
 * is rebuilt whenever model, tree or paths change, so nothing is left to do. */
void
SingleSiteSampler::reset(const EpiEvoModel &, vector<vector<Path> > &) {}

static uint64_t
seed_from(std::mt19937 &gen) {
  const uint64_t hi = gen(), lo = gen();
  return (hi << 32) | lo;
}

bool
SingleSiteSampler::Metropolis_Hastings_site(const EpiEvoModel &the_model, const TreeHelper &th,
                                            const size_t site_id,
                                            vector<vector<Path> > &paths, std::mt19937 &gen) {
  // see the header: one sweep per caller loop; the loop's return values add up to the accept count
  if (site_id != 1) return site_id <= last_sweep_accepts;
  configure(the_model, th);
  upload(paths, th);
  uint64_t n_acc = 0;
  const double t0 = now_s();
  check(epv_multi_sweep(ctx, seed_from(gen), sweep_counter++, &n_acc));
  phase_clock.device += now_s() - t0;
  download(paths);
  last_sweep_accepts = n_acc;
  return n_acc > 0;
}

void
SingleSiteSampler::run_mcmc(const EpiEvoModel &the_model, const TreeHelper &th,
                            vector<vector<Path> > &paths, std::mt19937 &gen,
                            vector<vector<double> > &J_all_sites,
                            vector<vector<double> > &D_all_sites, double &acceptance_rate) {
  configure(the_model, th);
  upload(paths, th);
  vector<double> J(th.n_nodes*8), D(th.n_nodes*8);
  const double t0 = now_s();
  check(epv_multi_run_mcmc(ctx, burn_in, batch, seed_from(gen), sweep_counter, J.data(), D.data(),
                           &acceptance_rate));
  phase_clock.device += now_s() - t0;
  sweep_counter += burn_in + batch;
  download(paths);
  J_all_sites.assign(th.n_nodes, vector<double>(8, 0.0));   // row 0 unused, as the reference
  D_all_sites.assign(th.n_nodes, vector<double>(8, 0.0));
  for (size_t b = 1; b < th.n_nodes; ++b)
    for (size_t t = 0; t < 8; ++t) {
      J_all_sites[b][t] = J[b*8 + t];
      D_all_sites[b][t] = D[b*8 + t];
    }
}

namespace epievo_b200 {

/* get_sufficient_statistics per branch (ParamEstimation.cpp:92-114) on the GPU.  The topology does not enter
 * the statistics (they are per branch); ss / pi only have to describe SOME tree in preorder. */
static void
suffstats_on_device(const vector<vector<Path> > &all_paths, const size_t n_nodes,
                    const vector<uint64_t> &ss, const vector<uint64_t> &pi,
                    vector<vector<double> > &J, vector<vector<double> > &D) {
  // a throw-away context: tree + paths only (no model needed for J/D)
  epv_ctx *c = nullptr;
  if (epv_create(&c, int_from_env("EPIEVO_B200_DEVICE", 0)) != 0 || !c)
    throw std::runtime_error("epievo_b200: no usable CUDA device (there is no CPU fallback)");
  struct Guard { epv_ctx *c; ~Guard() { epv_destroy(c); } } guard{c};
  auto ck = [&](int rc) { if (rc) throw std::runtime_error(std::string("epievo_b200: ") + epv_last_error(c)); };
  const size_t n_sites = all_paths.size();
  if (n_sites < 3 || n_nodes < 2) throw std::runtime_error("epievo_b200: need at least 3 sites and one branch");
  vector<double> br(n_nodes, 0.0);
  for (size_t b = 1; b < n_nodes; ++b) br[b] = all_paths[0][b].tot_time;   // rows carry the time
  ck(epv_set_tree(c, n_nodes, ss.data(), pi.data(), br.data()));
  vector<unsigned char> init;
  vector<uint64_t> off;
  vector<double> jmp, tot;
  flatten_paths(all_paths, n_nodes, init, off, jmp, tot);
  ck(epv_upload_paths(c, n_sites, n_nodes, init.data(), off.data(), jmp.data(), br.data()));
  vector<double> Jf(n_nodes*8), Df(n_nodes*8);
  ck(epv_suffstats(c, Jf.data(), Df.data()));
  J.assign(n_nodes, vector<double>(8, 0.0));
  D.assign(n_nodes, vector<double>(8, 0.0));
  for (size_t b = 1; b < n_nodes; ++b)
    for (size_t t = 0; t < 8; ++t) { J[b][t] = Jf[b*8 + t]; D[b][t] = Df[b*8 + t]; }
}

void
get_sufficient_statistics(const vector<vector<Path> > &all_paths, const TreeHelper &th,
                          vector<vector<double> > &J, vector<vector<double> > &D) {
  vector<uint64_t> ss(th.subtree_sizes.begin(), th.subtree_sizes.end());
  vector<uint64_t> pi(th.parent_ids.begin(), th.parent_ids.end());
  suffstats_on_device(all_paths, th.n_nodes, ss, pi, J, D);
}

void
get_sufficient_statistics(const vector<vector<Path> > &all_paths,
                          vector<vector<double> > &J, vector<vector<double> > &D) {
  if (all_paths.empty()) throw std::runtime_error("epievo_b200: no paths");
  const size_t n_nodes = all_paths[0].size();
  vector<uint64_t> ss(n_nodes, 1), pi(n_nodes, 0);       // a star: every node hangs off the root
  ss[0] = n_nodes;
  suffstats_on_device(all_paths, n_nodes, ss, pi, J, D);
}

}  // namespace epievo_b200
