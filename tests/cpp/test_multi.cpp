/* test_multi.cpp -- a plain C++ host program (no torch, no Python) that drives N contexts through the
 * C-ABI's multi-GPU handle (epv_multi_*, include/epievo_b200.h) and checks that the sharded chain
 * equals the single-context chain bit for bit: histories, J, accept rate; D to 1e-12.
 *   usage: test_multi [n_shards=2] [n_sites=3000]      (shards go to GPUs 0..k-1 round-robin over the
 *   devices present, so it runs on a one-GPU box too)
 * This is the host loop a maintainer of the reference gets through the drop-in
 * (EPIEVO_B200_GPUS, epievo_b200/host/SingleSiteSampler_b200.cpp). */
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include <cuda_runtime_api.h>

#include "epievo_b200.h"

static uint64_t lcg = 88172645463325252ull;
static double unif() { lcg ^= lcg << 13; lcg ^= lcg >> 7; lcg ^= lcg << 17; return (double)(lcg >> 11)*(1.0/9007199254740992.0); }

struct Paths {
  uint64_t n_sites, n_nodes;
  std::vector<uint8_t> init;
  std::vector<uint64_t> off;
  std::vector<double> jumps, tot;
};

/* valid random histories on the reference's test tree ((C:0.03,D:0.06)E:0.02,F:0.1)G */
static Paths make_paths(uint64_t n, const std::vector<uint64_t> &parent, const std::vector<double> &branch, double density) {
  Paths p; p.n_sites = n; p.n_nodes = parent.size();
  p.init.assign(p.n_nodes*n, 0); p.off.assign(p.n_nodes*n + 1, 0); p.tot = branch; p.tot[0] = 0.0;
  std::vector<std::vector<uint8_t>> end(p.n_nodes, std::vector<uint8_t>(n, 0));
  uint8_t x = 0;
  for (uint64_t s = 0; s < n; ++s) { if (unif() > 0.87) x ^= 1; end[0][s] = x; }
  for (uint64_t v = 1; v < p.n_nodes; ++v)
    for (uint64_t s = 0; s < n; ++s) {
      const uint64_t r = v*n + s;
      p.init[r] = end[parent[v]][s];
      p.off[r] = p.jumps.size();
      int k = 0;
      double t = 0.0;
      for (;;) {                                   // Poisson(density * T) sorted uniform times
        t += -std::log(1.0 - unif())/density;
        if (t >= branch[v]) break;
        p.jumps.push_back(t); ++k;
      }
      end[v][s] = p.init[r] ^ (k & 1);
    }
  p.off[p.n_nodes*n] = p.jumps.size();
  return p;
}

#define CK(m, call) do { if ((call) != 0) { std::fprintf(stderr, "FAILED %s: %s\n", #call, epv_multi_last_error(m)); return 1; } } while (0)

struct Result { std::vector<double> J, D; double acc; Paths out; };

static int run(int n_shards, const std::vector<int> &dev, const Paths &p, const std::vector<uint64_t> &subtree,
               const std::vector<uint64_t> &parent, const std::vector<double> &branch, const double *rates,
               const double *T, Result &res) {
  epv_multi *m = nullptr;
  if (epv_multi_create(&m, n_shards, dev.data()) != 0) { std::fprintf(stderr, "epv_multi_create failed\n"); return 1; }
  CK(m, epv_multi_set_tree(m, p.n_nodes, subtree.data(), parent.data(), branch.data()));
  CK(m, epv_multi_set_model(m, rates, T));
  CK(m, epv_multi_upload_paths(m, p.n_sites, p.n_nodes, p.init.data(), p.off.data(), p.jumps.data(), p.tot.data()));
  res.J.assign(p.n_nodes*8, 0.0); res.D.assign(p.n_nodes*8, 0.0);
  CK(m, epv_multi_run_mcmc(m, 3, 3, 4242, 7, res.J.data(), res.D.data(), &res.acc));
  uint64_t ns, nn, nj;
  CK(m, epv_multi_paths_size(m, &ns, &nn, &nj));
  res.out.n_sites = ns; res.out.n_nodes = nn;
  res.out.init.assign(ns*nn, 0); res.out.off.assign(ns*nn + 1, 0); res.out.jumps.assign(nj ? nj : 1, 0.0); res.out.tot.assign(nn, 0.0);
  CK(m, epv_multi_download_paths(m, res.out.init.data(), res.out.off.data(), res.out.jumps.data(), res.out.tot.data()));
  res.out.jumps.resize(nj);
  /* a rescale (the -b path's last step) must also agree */
  std::vector<double> nb(branch);
  for (size_t v = 1; v < nb.size(); ++v) nb[v] *= 1.25;
  CK(m, epv_multi_scale_jump_times(m, nb.data()));
  std::vector<double> J2(p.n_nodes*8), D2(p.n_nodes*8);
  CK(m, epv_multi_suffstats(m, J2.data(), D2.data()));
  for (size_t i = 0; i < J2.size(); ++i) { res.J.push_back(J2[i]); res.D.push_back(D2[i]); }
  epv_multi_destroy(m);
  return 0;
}

int main(int argc, char **argv) {
  const int n_shards = argc > 1 ? std::atoi(argv[1]) : 2;
  const uint64_t n = argc > 2 ? std::strtoull(argv[2], nullptr, 10) : 3000;
  int n_dev = 0;
  if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) { std::fprintf(stderr, "no CUDA device\n"); return 2; }
  const std::vector<uint64_t> subtree = {5, 3, 1, 1, 1}, parent = {0, 0, 1, 1, 0};
  const std::vector<double> branch = {0.0, 0.02, 0.03, 0.06, 0.1};
  double rates[8], T[4], Q[4], bl[4];
  epv_model_from_params(0.844912, 0.893359, -0.8, -1.8, 1, rates, T, Q, bl);     /* the reference's test/test.param */
  const Paths p = make_paths(n, parent, branch, 3.0);
  Result one, many;
  if (run(1, std::vector<int>(1, 0), p, subtree, parent, branch, rates, T, one)) return 1;
  std::vector<int> dev(n_shards);
  for (int r = 0; r < n_shards; ++r) dev[r] = r % n_dev;
  if (run(n_shards, dev, p, subtree, parent, branch, rates, T, many)) return 1;
  int bad = 0;
  if (one.acc != many.acc) { std::fprintf(stderr, "accept rate differs: %.17g vs %.17g\n", one.acc, many.acc); ++bad; }
  if (one.out.init != many.out.init || one.out.off != many.out.off || one.out.jumps != many.out.jumps) { std::fprintf(stderr, "histories differ\n"); ++bad; }
  for (size_t i = 0; i < one.J.size(); ++i) {
    if (one.J[i] != many.J[i]) { std::fprintf(stderr, "J[%zu] differs: %g vs %g\n", i, one.J[i], many.J[i]); ++bad; break; }
    if (std::fabs(one.D[i] - many.D[i]) > 1e-12*std::fabs(one.D[i])) { std::fprintf(stderr, "D[%zu] differs: %.17g vs %.17g\n", i, one.D[i], many.D[i]); ++bad; break; }
  }
  std::printf("%s: %d shards on %d device(s), %llu sites, acc %.6f, %zu jumps\n", bad ? "MISMATCH" : "ok", n_shards,
              n_shards < n_dev ? n_shards : n_dev, (unsigned long long)n, many.acc, many.out.jumps.size());
  return bad ? 1 : 0;
}
