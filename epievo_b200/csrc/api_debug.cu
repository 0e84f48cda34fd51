// api_debug.cu -- deterministic per-site intermediates and sampler-law probes used by
// the parity tests (they run the production device routines, not copies of them).
#include <cstring>
#include <vector>

#include "../../include/epievo_b200.h"
#include "context.h"

using namespace epv;

extern "C" {

int epv_site_debug(epv_ctx *c, uint64_t site, uint64_t cap, uint64_t *n_intervals,
                   uint64_t *seg_off, double *rate0, double *rate1, uint64_t *trip0,
                   uint64_t *trip1, double *len, double *p, double *q, double *root_p0,
                   double *proposal_logp) {
  if (!c) return 1;
  if (!c->n_sites || !c->have_model) return fail(c, "paths and model must be loaded");
  if (site < 1 || site + 1 >= c->n_sites) return fail(c, "site must be interior");
  cudaSetDevice(c->device);
  const int nn = c->n_nodes;
  // one device allocation, carved up
  const size_t n_u64 = 1 + (nn + 1) + 2*cap, n_dbl = 3*cap + 2*cap + 2*nn + 2;
  unsigned long long *d_u = nullptr; double *d_d = nullptr;
  EPV_CUDA(c, cudaMalloc(&d_u, n_u64*sizeof(unsigned long long)));
  EPV_CUDA(c, cudaMalloc(&d_d, n_dbl*sizeof(double)));
  EPV_CUDA(c, cudaMemsetAsync(d_u, 0, n_u64*sizeof(unsigned long long), c->stream));
  DebugOut o;
  o.n_intervals = d_u; o.seg_off = d_u + 1; o.trip0 = o.seg_off + nn + 1; o.trip1 = o.trip0 + cap;
  o.rate0 = d_d; o.rate1 = d_d + cap; o.len = d_d + 2*cap; o.p = d_d + 3*cap;
  o.q = o.p + 2*cap; o.root_p0 = o.q + 2*nn; o.logp = o.root_p0 + 1; o.cap = cap;
  launch_site_debug(c->mp, c->view(c->cur), c->d_state, c->W, c->tree(), site, o, c->stream);
  ++c->launches;
  std::vector<unsigned long long> hu(n_u64);
  std::vector<double> hd(n_dbl);
  EPV_CUDA(c, cudaMemcpyAsync(hu.data(), d_u, n_u64*sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaMemcpyAsync(hd.data(), d_d, n_dbl*sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  cudaFree(d_u); cudaFree(d_d);
  EPV_CUDA(c, cudaGetLastError());
  if (hu[0] == ~0ull) return fail(c, "epv_site_debug: interval capacity exceeded");
  const uint64_t k = hu[0];
  *n_intervals = k;
  for (int v = 0; v <= nn; ++v) seg_off[v] = hu[1 + v];
  for (uint64_t i = 0; i < k; ++i) {
    trip0[i] = hu[1 + nn + 1 + i]; trip1[i] = hu[1 + nn + 1 + cap + i];
    rate0[i] = hd[i]; rate1[i] = hd[cap + i]; len[i] = hd[2*cap + i];
    p[2*i] = hd[3*cap + 2*i]; p[2*i + 1] = hd[3*cap + 2*i + 1];
  }
  for (int v = 0; v < 2*nn; ++v) q[v] = hd[5*cap + v];
  *root_p0 = hd[5*cap + 2*nn];
  *proposal_logp = hd[5*cap + 2*nn + 1];
  return 0;
}

int epv_get_tri_llh(epv_ctx *c, double *tri_llh) {
  if (!c) return 1;
  if (!c->n_sites || !c->tri_valid) return fail(c, "call epv_reset first");
  cudaSetDevice(c->device);
  EPV_CUDA(c, cudaMemcpyAsync(tri_llh, c->d_tri, c->n_sites*sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  return 0;
}

int epv_end_cond_sample(epv_ctx *c, double rate0, double rate1, int start_state, int end_state,
                        double T, uint64_t seed, uint64_t n, uint64_t cap, uint32_t *n_jumps,
                        double *jumps) {
  if (!c) return 1;
  if (!(rate0 > 0 && rate1 > 0 && T > 0) || cap == 0) return fail(c, "bad sampler arguments");
  cudaSetDevice(c->device);
  uint32_t *d_n = nullptr; double *d_j = nullptr;
  EPV_CUDA(c, cudaMalloc(&d_n, n*sizeof(uint32_t)));
  EPV_CUDA(c, cudaMalloc(&d_j, n*cap*sizeof(double)));
  launch_end_cond(rate0, rate1, start_state, end_state, T, seed, n, cap, d_n, d_j, c->stream);
  ++c->launches;
  EPV_CUDA(c, cudaMemcpyAsync(n_jumps, d_n, n*sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaMemcpyAsync(jumps, d_j, n*cap*sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  cudaFree(d_n); cudaFree(d_j);
  EPV_CUDA(c, cudaGetLastError());
  return 0;
}

}  // extern "C"
