// api_multi.cu -- several GPUs behind ONE handle, in one host process (no MPI, no torch):
// contiguous site blocks, one context per GPU, 3-site halos, one boundary record per neighbour and
// colour pass moved by a peer copy over NVLink, J/D sums added on the host.  This is what the C++
// drop-in (host/SingleSiteSampler_b200.cpp, EPIEVO_B200_GPUS=N) drives, so the reference's programs
// use N GPUs unchanged (prog/epievo_est_params_histories.cpp:236-264 is the caller).
//
// Per colour pass and shard (epv_pass_begin / epv_pass_end, api_sampler.cu):
//   side stream : edge groups (halo-aware update) -> pack the boundary site's record from the staging
//                 blob -> peer copies into the neighbours' receive buffers -> event "sent"
//                 ... wait for the neighbours' "sent" -> unpack -> event "halo"
//   main stream : interior groups -> wait "edge done" -> CSR rebuild -> wait "halo"
// so the exchange (a few KB, latency only) hides behind the interior update.  Receive buffers are
// double-buffered by pass parity: a sender can be at most one pass ahead of a receiver's unpack.
// Philox is keyed by the GLOBAL site index: N shards reproduce the one-GPU chain bit for bit.
#include <algorithm>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/epievo_b200.h"
#include "context.h"

using namespace epv;

struct epv_multi {
  int n = 0;
  std::vector<epv_ctx *> ctx;
  std::vector<int> dev;
  std::vector<uint64_t> lo, hi;               // owned global sites [lo, hi) of every shard
  uint64_t n_global = 0;
  int n_nodes = 0;
  std::vector<double> h_branch;
  // halo records (device memory of the shard's GPU)
  uint64_t rec_bytes = 0;
  std::vector<void *> send_lo, send_hi;
  std::vector<void *> recv_lo[2], recv_hi[2];
  std::vector<cudaEvent_t> ev_sent;
  int parity = 0;
  void *stats_stage = nullptr;                // on shard 0's GPU: a peer's J/D accumulators in transit
  uint64_t stats_words = 0;
  std::string error;
};

namespace {

int mfail(epv_multi *m, const std::string &msg) { if (m) m->error = msg; return 1; }
int mctx(epv_multi *m, int r, int rc) {
  if (rc) m->error = std::string("shard ") + std::to_string(r) + ": " + epv_last_error(m->ctx[r]);
  return rc;
}
#define MCK(m, r, call) do { if (int _rc = mctx((m), (r), (call))) return _rc; } while (0)
#define MCUDA(m, call) do { cudaError_t _e = (call); if (_e != cudaSuccess) return mfail((m), std::string("CUDA error: ") + cudaGetErrorString(_e) + " in " #call); } while (0)

void free_records(epv_multi *m) {
  for (int r = 0; r < m->n; ++r) {
    cudaSetDevice(m->dev[r]);
    cudaFree(m->send_lo[r]); cudaFree(m->send_hi[r]);
    for (int p = 0; p < 2; ++p) { cudaFree(m->recv_lo[p][r]); cudaFree(m->recv_hi[p][r]); }
    m->send_lo[r] = m->send_hi[r] = nullptr;
    for (int p = 0; p < 2; ++p) m->recv_lo[p][r] = m->recv_hi[p][r] = nullptr;
  }
  m->rec_bytes = 0;
}

// Same halo-row capacity on every shard (the record layout depends on it) + record buffers.
int agree_halo_capacity(epv_multi *m, bool grow_check) {
  if (m->n == 1) return 0;
  uint64_t want = 0;
  for (int r = 0; r < m->n; ++r) {
    uint64_t cap = 0, need = 0;
    MCK(m, r, epv_halo_capacity(m->ctx[r], &cap, &need));
    want = std::max(want, grow_check ? std::max(cap, need) : cap);
  }
  for (int r = 0; r < m->n; ++r) MCK(m, r, epv_halo_set_capacity(m->ctx[r], want));
  uint64_t bytes = 0;
  MCK(m, 0, epv_halo_record_bytes(m->ctx[0], &bytes));
  if (bytes == m->rec_bytes) return 0;
  for (int r = 0; r < m->n; ++r) MCK(m, r, epv_synchronize(m->ctx[r]));
  free_records(m);
  for (int r = 0; r < m->n; ++r) {
    cudaSetDevice(m->dev[r]);
    MCUDA(m, cudaMalloc(&m->send_lo[r], bytes)); MCUDA(m, cudaMalloc(&m->send_hi[r], bytes));
    for (int p = 0; p < 2; ++p) { MCUDA(m, cudaMalloc(&m->recv_lo[p][r], bytes)); MCUDA(m, cudaMalloc(&m->recv_hi[p][r], bytes)); }
  }
  m->rec_bytes = bytes;
  return 0;
}

int copy_record(epv_multi *m, void *dst, int dst_r, const void *src, int src_r, cudaStream_t st) {
  if (m->dev[dst_r] == m->dev[src_r]) MCUDA(m, cudaMemcpyAsync(dst, src, m->rec_bytes, cudaMemcpyDeviceToDevice, st));
  else MCUDA(m, cudaMemcpyPeerAsync(dst, m->dev[dst_r], src, m->dev[src_r], m->rec_bytes, st));
  return 0;
}

// One colour pass over all shards (see the file header).
int multi_pass(epv_multi *m, int colour, uint64_t seed, uint64_t sweep) {
  const int n = m->n, p = m->parity;
  for (int r = 0; r < n; ++r)
    MCK(m, r, epv_pass_begin(m->ctx[r], colour, seed, sweep, r > 0 ? m->send_lo[r] : nullptr,
                             r + 1 < n ? m->send_hi[r] : nullptr));
  for (int r = 0; r < n; ++r) {
    cudaSetDevice(m->dev[r]);
    cudaStream_t st = m->ctx[r]->side_stream;
    if (r > 0) if (int rc = copy_record(m, m->recv_hi[p][r - 1], r - 1, m->send_lo[r], r, st)) return rc;
    if (r + 1 < n) if (int rc = copy_record(m, m->recv_lo[p][r + 1], r + 1, m->send_hi[r], r, st)) return rc;
    MCUDA(m, cudaEventRecord(m->ev_sent[r], st));
  }
  for (int r = 0; r < n; ++r) {
    cudaSetDevice(m->dev[r]);
    cudaStream_t st = m->ctx[r]->side_stream;
    if (r > 0) MCUDA(m, cudaStreamWaitEvent(st, m->ev_sent[r - 1], 0));
    if (r + 1 < n) MCUDA(m, cudaStreamWaitEvent(st, m->ev_sent[r + 1], 0));
    MCK(m, r, epv_pass_end(m->ctx[r], colour, r > 0 ? m->recv_lo[p][r] : nullptr,
                           r + 1 < n ? m->recv_hi[p][r] : nullptr));
  }
  m->parity ^= 1;
  return 0;
}

// J/D over all shards, exactly: every shard's integer accumulators are copied to shard 0's GPU and added
// there, then folded once -- the result does not depend on the number of shards.
int multi_fold_stats(epv_multi *m, double *J, double *D) {
  epv_ctx *c0 = m->ctx[0];
  uint64_t words = 0;
  MCK(m, 0, epv_stats_words(c0, &words));
  cudaSetDevice(m->dev[0]);
  if (words != m->stats_words) {
    cudaFree(m->stats_stage); m->stats_stage = nullptr;
    MCUDA(m, cudaMalloc(&m->stats_stage, words*8));
    m->stats_words = words;
  }
  for (int r = 1; r < m->n; ++r) {
    cudaSetDevice(m->dev[r]);
    MCUDA(m, cudaEventRecord(m->ev_sent[r], m->ctx[r]->stream));          // shard r's passes are enqueued
    cudaSetDevice(m->dev[0]);
    MCUDA(m, cudaStreamWaitEvent(c0->stream, m->ev_sent[r], 0));
    if (m->dev[r] == m->dev[0]) MCUDA(m, cudaMemcpyAsync(m->stats_stage, m->ctx[r]->d_ss, words*8, cudaMemcpyDeviceToDevice, c0->stream));
    else MCUDA(m, cudaMemcpyPeerAsync(m->stats_stage, m->dev[0], m->ctx[r]->d_ss, m->dev[r], words*8, c0->stream));
    MCK(m, 0, epv_stats_add(c0, m->stats_stage));
  }
  return mctx(m, 0, epv_stats_fold(c0, nullptr, J, D));
}

int multi_sweep(epv_multi *m, uint64_t seed, uint64_t sweep) {
  if (m->n == 1) return mctx(m, 0, epv_sweep(m->ctx[0], seed, sweep, nullptr));
  for (int colour = 0; colour < 3; ++colour)
    if (int rc = multi_pass(m, colour, seed, sweep)) return rc;
  return agree_halo_capacity(m, true);      // boundary rows that grew: regrow before they overflow
}

// node-major CSR of global sites [a, b) cut out of the global arrays
void slice_paths(uint64_t n_sites, uint64_t n_nodes, const uint8_t *init, const uint64_t *off, const double *jumps,
                 uint64_t a, uint64_t b, std::vector<uint8_t> &l_init, std::vector<uint64_t> &l_off,
                 std::vector<double> &l_jmp) {
  const uint64_t w = b - a;
  l_init.resize(n_nodes*w); l_off.resize(n_nodes*w + 1);
  uint64_t total = 0;
  for (uint64_t v = 0; v < n_nodes; ++v) total += off[v*n_sites + b] - off[v*n_sites + a];
  l_jmp.resize(total ? total : 1);
  uint64_t k = 0;
  for (uint64_t v = 0; v < n_nodes; ++v) {
    const uint64_t g0 = v*n_sites + a, base = off[g0];
    std::memcpy(&l_init[v*w], init + g0, w);
    for (uint64_t s = 0; s < w; ++s) l_off[v*w + s] = k + (off[g0 + s] - base);
    const uint64_t cnt = off[g0 + w] - base;
    if (cnt) std::memcpy(&l_jmp[k], jumps + base, cnt*sizeof(double));
    k += cnt;
  }
  l_off[n_nodes*w] = k;
}

}  // namespace

extern "C" {

int epv_multi_create(epv_multi **out, int n_shards, const int *devices) {
  if (!out) return 1;
  *out = nullptr;
  if (n_shards < 1 || n_shards > 64) return 1;
  epv_multi *m = new epv_multi;
  m->n = n_shards;
  m->ctx.assign(n_shards, nullptr); m->dev.resize(n_shards);
  m->send_lo.assign(n_shards, nullptr); m->send_hi.assign(n_shards, nullptr);
  for (int p = 0; p < 2; ++p) { m->recv_lo[p].assign(n_shards, nullptr); m->recv_hi[p].assign(n_shards, nullptr); }
  m->ev_sent.assign(n_shards, nullptr);
  for (int r = 0; r < n_shards; ++r) {
    m->dev[r] = devices ? devices[r] : r;
    if (epv_create(&m->ctx[r], m->dev[r]) != 0) { epv_multi_destroy(m); return 1; }
    cudaSetDevice(m->dev[r]);
    cudaEventCreateWithFlags(&m->ev_sent[r], cudaEventDisableTiming);
  }
  // peer access between ring neighbours (records move by cudaMemcpyPeerAsync; without peer access the
  // runtime stages them through the host, still correct)
  for (int r = 0; r + 1 < n_shards; ++r) {
    const int a = m->dev[r], b = m->dev[r + 1];
    if (a == b) continue;
    int ok = 0;
    if (cudaDeviceCanAccessPeer(&ok, a, b) == cudaSuccess && ok) { cudaSetDevice(a); cudaDeviceEnablePeerAccess(b, 0); }
    if (cudaDeviceCanAccessPeer(&ok, b, a) == cudaSuccess && ok) { cudaSetDevice(b); cudaDeviceEnablePeerAccess(a, 0); }
    cudaGetLastError();                    // "already enabled" is fine
  }
  *out = m;
  return 0;
}

void epv_multi_destroy(epv_multi *m) {
  if (!m) return;
  for (int r = 0; r < m->n; ++r) if (m->ctx[r]) epv_synchronize(m->ctx[r]);
  free_records(m);
  if (m->stats_stage) { cudaSetDevice(m->dev[0]); cudaFree(m->stats_stage); }
  for (int r = 0; r < m->n; ++r) {
    if (m->ev_sent[r]) { cudaSetDevice(m->dev[r]); cudaEventDestroy(m->ev_sent[r]); }
    epv_destroy(m->ctx[r]);
  }
  delete m;
}

const char *epv_multi_last_error(const epv_multi *m) { return m ? m->error.c_str() : "null handle"; }
int epv_multi_shards(const epv_multi *m) { return m ? m->n : 0; }
epv_ctx *epv_multi_context(epv_multi *m, int shard) { return m && shard >= 0 && shard < m->n ? m->ctx[shard] : nullptr; }

int epv_multi_set_tree(epv_multi *m, uint64_t n_nodes, const uint64_t *subtree_sizes, const uint64_t *parent_ids,
                       const double *branches) {
  if (!m) return 1;
  for (int r = 0; r < m->n; ++r) MCK(m, r, epv_set_tree(m->ctx[r], n_nodes, subtree_sizes, parent_ids, branches));
  m->n_nodes = (int)n_nodes;
  m->h_branch.assign(branches, branches + n_nodes);
  return 0;
}

int epv_multi_set_model(epv_multi *m, const double *rates, const double *T) {
  if (!m) return 1;
  for (int r = 0; r < m->n; ++r) MCK(m, r, epv_set_model(m->ctx[r], rates, T));
  return 0;
}

/* Contiguous blocks whose starts are multiples of 3 (local and global colours agree). */
int epv_multi_partition(uint64_t n_global, int n_shards, uint64_t *begin, uint64_t *end) {
  const uint64_t n3 = n_global/3;
  for (int r = 0; r < n_shards; ++r) {
    begin[r] = 3*((n3*(uint64_t)r)/(uint64_t)n_shards);
    if (r) end[r - 1] = begin[r];
  }
  end[n_shards - 1] = n_global;
  for (int r = 0; r < n_shards; ++r) if (end[r] - begin[r] < 2*HALO_W) return 1;
  return 0;
}

int epv_multi_upload_paths(epv_multi *m, uint64_t n_sites, uint64_t n_nodes, const uint8_t *init_state,
                           const uint64_t *row_off, const double *jumps, const double *tot_time) {
  if (!m) return 1;
  const int n = m->n;
  if (n == 1) {
    m->n_global = n_sites; m->lo.assign(1, 0); m->hi.assign(1, n_sites);
    return mctx(m, 0, epv_upload_paths(m->ctx[0], n_sites, n_nodes, init_state, row_off, jumps, tot_time));
  }
  std::vector<uint64_t> lo(n), hi(n);
  if (epv_multi_partition(n_sites, n, lo.data(), hi.data())) return mfail(m, "too few sites for this many shards (each needs 6)");
  if (n_sites != m->n_global || lo != m->lo) {
    // (re)shard: epv_set_shard only before paths are loaded
    for (int r = 0; r < n; ++r)
      if (m->ctx[r]->n_sites) { cudaSetDevice(m->dev[r]); cudaStreamSynchronize(m->ctx[r]->stream); free_paths(m->ctx[r]); }
    for (int r = 0; r < n; ++r)
      MCK(m, r, epv_set_shard(m->ctx[r], n_sites, lo[r], hi[r], r > 0 ? HALO_W : 0, r + 1 < n ? HALO_W : 0));
    m->n_global = n_sites; m->lo = lo; m->hi = hi;
  }
  // slice on host threads (one per shard), then upload (each upload is asynchronous up to its last sync)
  std::vector<std::vector<uint8_t>> l_init(n);
  std::vector<std::vector<uint64_t>> l_off(n);
  std::vector<std::vector<double>> l_jmp(n);
  std::vector<std::thread> th;
  for (int r = 0; r < n; ++r)
    th.emplace_back([&, r] {
      slice_paths(n_sites, n_nodes, init_state, row_off, jumps, lo[r] - (r > 0 ? HALO_W : 0),
                  hi[r] + (r + 1 < n ? HALO_W : 0), l_init[r], l_off[r], l_jmp[r]);
    });
  for (auto &t : th) t.join();
  for (int r = 0; r < n; ++r) {
    const uint64_t w = hi[r] - lo[r] + (r > 0 ? HALO_W : 0) + (r + 1 < n ? HALO_W : 0);
    MCK(m, r, epv_upload_paths(m->ctx[r], w, n_nodes, l_init[r].data(), l_off[r].data(), l_jmp[r].data(), tot_time));
  }
  m->parity = 0;
  return agree_halo_capacity(m, false);
}

int epv_multi_paths_size(epv_multi *m, uint64_t *n_sites, uint64_t *n_nodes, uint64_t *n_jumps) {
  if (!m) return 1;
  if (!m->n_global) return mfail(m, "no paths loaded");
  uint64_t total = 0;
  std::vector<uint64_t> per((size_t)m->n_nodes);
  for (int r = 0; r < m->n; ++r) {
    MCK(m, r, epv_jump_counts(m->ctx[r], per.data()));          // owned rows only
    for (int v = 0; v < m->n_nodes; ++v) total += per[v];
  }
  if (n_sites) *n_sites = m->n_global;
  if (n_nodes) *n_nodes = (uint64_t)m->n_nodes;
  if (n_jumps) *n_jumps = total;
  return 0;
}

int epv_multi_download_paths(epv_multi *m, uint8_t *init_state, uint64_t *row_off, double *jumps, double *tot_time) {
  if (!m) return 1;
  if (!m->n_global) return mfail(m, "no paths loaded");
  const int n = m->n;
  if (n == 1) return mctx(m, 0, epv_download_paths(m->ctx[0], init_state, row_off, jumps, tot_time));
  const uint64_t N = m->n_global, nn = (uint64_t)m->n_nodes;
  std::vector<std::vector<uint8_t>> l_init(n);
  std::vector<std::vector<uint64_t>> l_off(n);
  std::vector<std::vector<double>> l_jmp(n);
  std::vector<uint64_t> width(n);
  for (int r = 0; r < n; ++r) {
    uint64_t ns = 0, nd = 0, nj = 0;
    MCK(m, r, epv_paths_size(m->ctx[r], &ns, &nd, &nj));
    width[r] = ns;
    l_init[r].resize(ns*nd); l_off[r].resize(ns*nd + 1); l_jmp[r].resize(nj ? nj : 1);
    MCK(m, r, epv_download_paths(m->ctx[r], l_init[r].data(), l_off[r].data(), l_jmp[r].data(), tot_time));
  }
  uint64_t k = 0;
  for (uint64_t v = 0; v < nn; ++v)
    for (int r = 0; r < n; ++r) {
      const uint64_t hl = r > 0 ? HALO_W : 0, w = width[r], own = m->hi[r] - m->lo[r];
      const uint64_t l0 = v*w + hl, base = l_off[r][l0];
      std::memcpy(init_state + v*N + m->lo[r], &l_init[r][l0], own);
      for (uint64_t s = 0; s < own; ++s) row_off[v*N + m->lo[r] + s] = k + (l_off[r][l0 + s] - base);
      const uint64_t cnt = l_off[r][l0 + own] - base;
      if (cnt) std::memcpy(jumps + k, &l_jmp[r][base], cnt*sizeof(double));
      k += cnt;
    }
  row_off[nn*N] = k;
  return 0;
}

int epv_multi_reset(epv_multi *m) {
  if (!m) return 1;
  for (int r = 0; r < m->n; ++r) MCK(m, r, epv_reset(m->ctx[r]));
  return 0;
}

int epv_multi_sweep(epv_multi *m, uint64_t seed, uint64_t sweep_idx, uint64_t *n_accepted) {
  if (!m) return 1;
  if (m->n == 1) return mctx(m, 0, epv_sweep(m->ctx[0], seed, sweep_idx, n_accepted));
  for (int r = 0; r < m->n; ++r) MCK(m, r, epv_accept_count(m->ctx[r], 1, nullptr));
  if (int rc = multi_sweep(m, seed, sweep_idx)) return rc;
  uint64_t tot = 0;
  for (int r = 0; r < m->n; ++r) { uint64_t a = 0; MCK(m, r, epv_accept_count(m->ctx[r], 1, &a)); tot += a; }
  if (n_accepted) *n_accepted = tot;
  return 0;
}

int epv_multi_suffstats(epv_multi *m, double *J, double *D) {
  if (!m) return 1;
  if (m->n == 1) return mctx(m, 0, epv_suffstats(m->ctx[0], J, D));
  for (int r = 0; r < m->n; ++r) { MCK(m, r, epv_stats_begin(m->ctx[r])); MCK(m, r, epv_stats_accumulate(m->ctx[r])); }
  return multi_fold_stats(m, J, D);
}

int epv_multi_run_mcmc(epv_multi *m, uint64_t burn_in, uint64_t batch, uint64_t seed, uint64_t first_sweep,
                       double *J, double *D, double *acc_rate) {
  if (!m) return 1;
  if (!m->n_global) return mfail(m, "no paths loaded");
  if (m->n == 1) return mctx(m, 0, epv_run_mcmc(m->ctx[0], burn_in, batch, seed, first_sweep, J, D, acc_rate));
  const size_t k = (size_t)m->n_nodes*8;
  uint64_t sweep = first_sweep;
  for (uint64_t i = 0; i < burn_in; ++i, ++sweep)            // SingleSiteSampler.cpp:560-561
    if (int rc = multi_sweep(m, seed, sweep)) return rc;
  for (int r = 0; r < m->n; ++r) MCK(m, r, epv_accept_count(m->ctx[r], 1, nullptr));
  for (int r = 0; r < m->n; ++r) MCK(m, r, epv_stats_begin(m->ctx[r]));
  for (uint64_t i = 0; i < batch; ++i, ++sweep) {            // :576-587
    if (int rc = multi_sweep(m, seed, sweep)) return rc;
    for (int r = 0; r < m->n; ++r) MCK(m, r, epv_stats_accumulate(m->ctx[r]));     // enqueued only
  }
  if (int rc = multi_fold_stats(m, J, D)) return rc;
  uint64_t acc = 0;
  for (int r = 0; r < m->n; ++r) { uint64_t a = 0; MCK(m, r, epv_accept_count(m->ctx[r], 1, &a)); acc += a; }
  for (size_t x = 0; x < k; ++x) { J[x] /= (double)batch; D[x] /= (double)batch; }      // :590-594
  if (acc_rate) *acc_rate = (double)acc/((double)batch*(double)(m->n_global - 2));      // :597
  return 0;
}

int epv_multi_scale_jump_times(epv_multi *m, const double *new_branches) {
  if (!m) return 1;
  for (int r = 0; r < m->n; ++r) MCK(m, r, epv_scale_jump_times(m->ctx[r], new_branches));
  m->h_branch.assign(new_branches, new_branches + m->n_nodes);
  return 0;
}

int epv_multi_mcem_iteration(epv_multi *m, uint64_t burn_in, uint64_t batch, uint64_t seed, uint64_t first_sweep,
                             int optimize_branches, double param_tol, double *rates, double *T, double *branches,
                             double *baseline, double *acc_rate, double *llh) {
  if (!m) return 1;
  if (m->n == 1)
    return mctx(m, 0, epv_mcem_iteration(m->ctx[0], burn_in, batch, seed, first_sweep, optimize_branches, param_tol,
                                         rates, T, branches, baseline, acc_rate, llh));
  if (int rc = epv_multi_set_model(m, rates, T)) return rc;
  std::vector<double> J((size_t)m->n_nodes*8), D((size_t)m->n_nodes*8);
  if (int rc = epv_multi_run_mcmc(m, burn_in, batch, seed, first_sweep, J.data(), D.data(), acc_rate)) return rc;
  double bl[4];
  if (!optimize_branches) {                                  // epievo_est_params_histories.cpp:254-257
    *llh = epv_estimate_rates((uint64_t)m->n_nodes, J.data(), D.data(), param_tol, rates, T, bl);
    epv_set_one_change(rates, (uint64_t)m->n_nodes, branches);
  }
  else                                                       // :259-262
    *llh = epv_estimate_rates_and_branches((uint64_t)m->n_nodes, J.data(), D.data(), param_tol, rates, branches, T, bl);
  if (baseline) for (int i = 0; i < 4; ++i) baseline[i] = bl[i];
  return epv_multi_scale_jump_times(m, branches);            // :264
}

uint64_t epv_multi_launch_count(const epv_multi *m) {
  uint64_t t = 0;
  if (m) for (int r = 0; r < m->n; ++r) t += epv_launch_count(m->ctx[r]);
  return t;
}

}  // extern "C"
