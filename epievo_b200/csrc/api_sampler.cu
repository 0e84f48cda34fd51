// api_sampler.cu -- reset / sweep / run_mcmc / sufficient statistics / rescale.
#include <cstring>
#include <vector>

#include "../../include/epievo_b200.h"
#include "context.h"

using namespace epv;

namespace epv {
int read_branch_totals(epv_ctx *c, std::vector<uint64_t> &tot);
int maybe_regrow(epv_ctx *c, const std::vector<uint64_t> &tot);
int regrow_to(epv_ctx *c, const std::vector<uint64_t> &cap, const std::vector<uint64_t> &tot, bool keep_stage);
int alloc_stage(epv_ctx *c);
}

namespace {

int ready(epv_ctx *c) {
  if (!c) return 1;
  if (!c->n_sites) return fail(c, "no paths loaded");
  if (!c->have_model) return fail(c, "epv_set_model has not been called");
  cudaSetDevice(c->device);
  return 0;
}

// Owned sites that have both neighbours globally: the J/D and tri_llh range.
void interior(const epv_ctx *c, uint64_t &lo, uint64_t &hi) {
  lo = c->upd_lo(); hi = c->upd_hi();
  const int64_t g0 = c->global0();
  if (g0 + (int64_t)lo < 1) lo = (uint64_t)(1 - g0);
  if (g0 + (int64_t)hi > (int64_t)c->n_global - 1) hi = (uint64_t)((int64_t)c->n_global - 1 - g0);
}

// Resolve the pending (start, stop) event pairs of the colour-pass launches.
void drain_timing(epv_ctx *c) {
  if (!c->ev_pending) return;
  cudaEventSynchronize(c->ev_ring[3*c->ev_pending - 1]);
  for (int k = 0; k < c->ev_pending; ++k) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, c->ev_ring[3*k], c->ev_ring[3*k + 1]) == cudaSuccess) c->mh_ms += ms;
    if (cudaEventElapsedTime(&ms, c->ev_ring[3*k + 1], c->ev_ring[3*k + 2]) == cudaSuccess) c->rebuild_ms += ms;
  }
  c->ev_pending = 0;
}

int do_reset(epv_ctx *c) {
  if (int r = alloc_stage(c)) return r;        // the staging is sized from the model, which may be newer than the paths
  launch_tri_llh(c->mp, c->view(c->cur), c->d_state, c->W, c->tree(), c->n_sites, c->d_tri, c->stream);
  ++c->launches;
  c->tri_valid = true;
  return 0;
}

struct IndepLaunch { const IndepTab *tab; double r0, r1; };

// split: the halo-aware edge groups run FIRST, on the context's side stream, so that the boundary
// site's record can be packed (from the staging blob) and sent while the interior groups and the CSR
// rebuild run on the main stream (epv_pass_begin / epv_pass_end).
int one_pass(epv_ctx *c, int colour, uint64_t seed, uint32_t sweep, const IndepLaunch *indep = nullptr,
             bool split = false) {
  if (++c->epoch >= (1u << 30)) {
    EPV_CUDA(c, cudaMemsetAsync(c->d_desc, 0, (size_t)(c->n_nodes - 1)*c->n_tiles*sizeof(unsigned long long), c->stream));
    c->epoch = 1;
  }
  PassArgs a;
  a.mp = c->mp;
  a.in = c->view(c->cur);
  a.off_out = c->d_off[c->cur ^ 1]; a.jmp_out = c->d_jmp[c->cur ^ 1];
  a.jcap = c->d_jcap; a.jneed = c->d_jneed;
  a.state = c->d_state; a.W = c->W; a.tree = c->tree(); a.tri_llh = c->d_tri;
  a.n_sites = c->n_sites; a.upd_lo = c->upd_lo(); a.upd_hi = c->upd_hi();
  a.global0 = c->global0(); a.n_global = c->n_global;
  a.colour = colour; a.seed = seed; a.sweep = sweep;
  a.q_slots = 0;
  for (int v = 1; v < c->n_nodes; ++v) a.q_slots += c->h_subtree[v] > 1;
  // sparse rows (0.05 jumps each on the 8-leaf tree): few lanes have jump times to draw, so
  // the warp pools them (5.5 vs 6.5 ms per pass); dense rows (1 jump each on the pairwise
  // branch): every lane is busy anyway and pooling only adds traffic (0.44 vs 0.39 ms)
  a.coop = c->jumps_per_row < 0.5;
  a.desc = c->d_desc; a.n_tiles = c->n_tiles; a.epoch = c->epoch;
  a.ticket = c->d_ticket; a.ticket_base = c->ticket_base;
  a.scratch = c->d_scratch; a.scratch_ctr = c->d_scratch_ctr;
  a.scratch_slots = c->scratch_slots; a.scratch_cap = c->scratch_cap;
  a.err = c->d_err; a.n_accept = c->d_nacc;
  a.n_groups = c->n_groups; a.stage_off = c->d_stg_off; a.stage = c->d_stg;
  a.stage_ctr = c->d_stg_ctr; a.stage_cap = c->stg_cap;
  a.ovf_list = c->d_ovf_list; a.list_mode = 0; a.reverse = 0;
  EPV_CUDA(c, cudaMemsetAsync(c->d_ovf_list, 0, sizeof(uint32_t), c->stream));
  EPV_CUDA(c, cudaMemsetAsync(c->d_scratch_ctr, 0, sizeof(uint32_t), c->stream));
  EPV_CUDA(c, cudaMemsetAsync(c->d_stg_ctr, 0, sizeof(unsigned long long), c->stream));
  if (c->timing) {
    if (c->ev_pending == epv_ctx::EV_RING) drain_timing(c);
    cudaEventRecord(c->ev_ring[3*c->ev_pending], c->stream);
  }
  int n_upd = 1;
  if (indep) launch_indep_sample(a, indep->tab, indep->r0, indep->r1, c->stream);
  else if (!split) n_upd = launch_mh_pass(a, c->stream);        // 1, or 2 for a sharded context
  else {
    EPV_CUDA(c, cudaEventRecord(c->ev_begin, c->stream));        // counters reset, previous rebuild done
    EPV_CUDA(c, cudaStreamWaitEvent(c->side_stream, c->ev_begin, 0));
    n_upd = launch_mh_edge(a, c->side_stream);
    EPV_CUDA(c, cudaEventRecord(c->ev_edge, c->side_stream));
    n_upd += launch_mh_interior(a, c->stream);
  }
  if (c->timing) {
    cudaEventRecord(c->ev_ring[3*c->ev_pending + 1], c->stream);
  }
  if (split) EPV_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_edge, 0));   // the rebuild reads every group's staging
  launch_csr_rebuild(a, c->stream);
  if (c->timing) {
    cudaEventRecord(c->ev_ring[3*c->ev_pending + 2], c->stream);
    ++c->ev_pending;
  }
  c->launches += n_upd + 1;
  if (!indep) ++c->mh_launches;
  c->ticket_base += c->n_tiles;
  // Thin headroom (or the site-independent resampling, whose rows can grow several-fold): look at the error
  // word now.  A rebuild that overflowed a branch wrote nothing usable, but its inputs -- the CSR the pass
  // read and the staging blob -- are intact and it recorded the totals it needs: regrow, rerun the rebuild.
  if (c->guard || indep) {
    for (int attempt = 0; attempt < 4; ++attempt) {
      unsigned flags = 0;
      EPV_CUDA(c, cudaMemcpyAsync(c->h_mail, c->d_err, sizeof(unsigned), cudaMemcpyDeviceToHost, c->stream));
      EPV_CUDA(c, cudaStreamSynchronize(c->stream));
      flags = *reinterpret_cast<unsigned *>(c->h_mail);
      if (!(flags & ERR_JUMP_CAPACITY)) break;
      const int n_br = c->n_nodes - 1;
      std::vector<unsigned long long> need(n_br);
      EPV_CUDA(c, cudaMemcpyAsync(need.data(), c->d_jneed, n_br*sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
      EPV_CUDA(c, cudaStreamSynchronize(c->stream));
      std::vector<uint64_t> tot;                                 // fill of the INPUT CSR (c->cur is still the input)
      if (int r = read_branch_totals(c, tot)) return r;
      std::vector<uint64_t> cap(n_br);
      for (int bi = 0; bi < n_br; ++bi) {
        if (need[bi] > 0xfffffff0ull) return fail(c, "a branch holds more than 2^32 jumps on one GPU; use more shards");
        const uint64_t want = need[bi] ? 3*need[bi] + c->n_own()/16 + 4096 : 0;
        cap[bi] = std::max<uint64_t>(c->h_jcap[bi], (want + 15) & ~15ull);
      }
      if (int r = regrow_to(c, cap, tot, true)) return r;
      flags &= ~ERR_JUMP_CAPACITY;
      EPV_CUDA(c, cudaMemcpyAsync(c->d_err, &flags, sizeof(unsigned), cudaMemcpyHostToDevice, c->stream));
      EPV_CUDA(c, cudaMemsetAsync(c->d_jneed, 0, n_br*sizeof(unsigned long long), c->stream));
      EPV_CUDA(c, cudaStreamSynchronize(c->stream));
      ++c->epoch;
      a.in = c->view(c->cur);
      a.jmp_out = c->d_jmp[c->cur ^ 1];
      a.epoch = c->epoch; a.ticket_base = c->ticket_base;
      launch_csr_rebuild(a, c->stream);
      c->ticket_base += c->n_tiles;
      ++c->launches; ++c->recoveries;
    }
  }
  c->cur ^= 1;
  c->stats_valid = false;
  return 0;
}

// One sweep = three colour passes.  Single-GPU shards need no exchange in between.
int one_sweep(epv_ctx *c, uint64_t seed, uint64_t sweep_idx) {
  for (int colour = 0; colour < 3; ++colour)
    if (int r = one_pass(c, colour, seed, (uint32_t)sweep_idx)) return r;
  return 0;
}

// After a sweep: surface device-side failures and keep capacity headroom.  One strided copy of the
// branch totals + the error word, one synchronisation.
int housekeeping(epv_ctx *c) {
  const int n_br = c->n_nodes - 1;
  uint32_t *mail = reinterpret_cast<uint32_t *>(c->h_mail);
  EPV_CUDA(c, cudaMemcpy2DAsync(mail + 1, sizeof(uint32_t), c->d_off[c->cur] + c->n_own(), c->off_pitch*sizeof(uint32_t),
                                sizeof(uint32_t), (size_t)n_br, cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaMemcpyAsync(mail, c->d_err, sizeof(unsigned), cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  if (c->halo_lo || c->halo_hi) EPV_CUDA(c, cudaStreamSynchronize(c->side_stream));
  if (mail[0]) return check_device_flags(c);
  std::vector<uint64_t> tot(n_br);
  uint64_t sum = 0;
  for (int bi = 0; bi < n_br; ++bi) sum += tot[bi] = mail[1 + bi];
  c->jumps_per_row = (double)sum/((double)n_br*(double)c->n_own());
  return maybe_regrow(c, tot);
}

// The same check read a sweep LATE, without stalling the device: the totals and the error word of sweep k are
// copied asynchronously and looked at after sweep k+1 was enqueued-and-finished (the event is long complete),
// so the GPU never waits for the host between sweeps (a synchronisation costs ~30 us of idle device per
// sweep from C++ and ~0.2 ms from a Python driver: 1 % of a headline step, 8 % of a strong-scaling step at 8
// GPUs).  Totals move by a fraction of a per cent per sweep and the regrow rule fires at half of the
// capacity, so one sweep of lag is harmless; when the late totals do ask for more room -- or headroom is thin
// (guard) -- the synchronous path runs with fresh totals.  housekeeping_take must follow the last sweep of a
// run (every entry point that returns results calls it).
int housekeeping_lagged(epv_ctx *c) {
  if (c->guard || c->n_nodes - 1 > 4000) { c->hk_pending = false; return housekeeping(c); }
  if (int r = housekeeping_take(c)) return r;
  const int n_br = c->n_nodes - 1;
  EPV_CUDA(c, cudaMemcpy2DAsync(c->h_hk + 1, sizeof(uint32_t), c->d_off[c->cur] + c->n_own(), c->off_pitch*sizeof(uint32_t),
                                sizeof(uint32_t), (size_t)n_br, cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaMemcpyAsync(c->h_hk, c->d_err, sizeof(unsigned), cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaEventRecord(c->ev_hk, c->stream));
  c->hk_pending = true;
  return 0;
}

int read_accepts(epv_ctx *c, uint64_t *n) {
  if (int r = housekeeping_take(c)) return r;
  EPV_CUDA(c, cudaMemcpyAsync(c->h_mail, c->d_nacc, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  *n = c->h_mail[0];
  return 0;
}

int do_suffstats(epv_ctx *c, bool clear = true, bool fold = true) {
  uint64_t lo, hi; interior(c, lo, hi);
  c->launches += launch_suffstats(c->view(c->cur), c->d_state, c->W, c->tree(), c->n_sites, lo, hi,
                                  c->d_J, c->d_D, c->d_ss, c->stream, c->sm_count, clear, fold);
  c->stats_valid = fold;
  return 0;
}

}  // namespace

namespace epv {

int housekeeping_take(epv_ctx *c) {
  if (!c->hk_pending) return 0;
  c->hk_pending = false;
  EPV_CUDA(c, cudaEventSynchronize(c->ev_hk));
  if (c->h_hk[0]) return check_device_flags(c);
  const int n_br = c->n_nodes - 1;
  std::vector<uint64_t> tot(n_br);
  uint64_t sum = 0;
  for (int bi = 0; bi < n_br; ++bi) sum += tot[bi] = c->h_hk[1 + bi];
  c->jumps_per_row = (double)sum/((double)n_br*(double)c->n_own());
  for (int bi = 0; bi < n_br; ++bi)
    if (2*tot[bi] > c->h_jcap[bi] || tot[bi] > 0xfffffff0ull) return housekeeping(c);
  return maybe_regrow(c, tot);          // (no growth: refreshes the guard flag and the staging size)
}

}  // namespace epv

extern "C" {

int epv_reset(epv_ctx *c) {
  if (int r = ready(c)) return r;
  if (int r = do_reset(c)) return r;
  EPV_CUDA(c, cudaGetLastError());
  return 0;
}

int epv_sweep(epv_ctx *c, uint64_t seed, uint64_t sweep_idx, uint64_t *n_accepted) {
  if (int r = ready(c)) return r;
  if (!c->tri_valid) if (int r = do_reset(c)) return r;
  EPV_CUDA(c, cudaMemsetAsync(c->d_nacc, 0, sizeof(unsigned long long), c->stream));
  if (int r = one_sweep(c, seed, sweep_idx)) return r;
  if (int r = housekeeping(c)) return r;
  uint64_t n = 0;
  if (int r = read_accepts(c, &n)) return r;
  if (n_accepted) *n_accepted = n;
  EPV_CUDA(c, cudaGetLastError());
  return 0;
}

int epv_run_mcmc_raw(epv_ctx *c, uint64_t burn_in, uint64_t batch, uint64_t seed,
                     uint64_t first_sweep, double *J_sum, double *D_sum, uint64_t *n_accepted) {
  if (int r = ready(c)) return r;
  if (c->have_shard && (c->halo_lo || c->halo_hi))
    return fail(c, "epv_run_mcmc needs a halo exchange between colour passes on a sharded "
                   "context; drive epv_pass/epv_halo_* from the multi-GPU host loop instead");
  if (!c->tri_valid) if (int r = do_reset(c)) return r;
  const size_t nb = (size_t)c->n_nodes*8*sizeof(double);
  uint64_t sweep = first_sweep;
  for (uint64_t i = 0; i < burn_in; ++i, ++sweep) {          // SingleSiteSampler.cpp:560-561
    if (int r = one_sweep(c, seed, sweep)) return r;
    if (int r = housekeeping_lagged(c)) return r;
  }
  EPV_CUDA(c, cudaMemsetAsync(c->d_nacc, 0, sizeof(unsigned long long), c->stream));
  // the batch's J/D passes add into ONE set of integer accumulators, folded once: the sums over the
  // batch are exact, and the same whatever the number of shards (epv_multi_run_mcmc adds the shards'
  // accumulators before the fold)
  EPV_CUDA(c, cudaMemsetAsync(c->d_ss, 0, suffstats_scratch_bytes(c->n_nodes, c->sm_count), c->stream));
  for (uint64_t i = 0; i < batch; ++i, ++sweep) {            // :576-587
    if (int r = one_sweep(c, seed, sweep)) return r;
    if (int r = housekeeping_lagged(c)) return r;
    if (int r = do_suffstats(c, false, false)) return r;
  }
  launch_suffstats_fold(c->d_ss, c->tree(), c->d_Jsum, c->d_Dsum, c->stream);
  ++c->launches;
  EPV_CUDA(c, cudaMemcpyAsync(J_sum, c->d_Jsum, nb, cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaMemcpyAsync(D_sum, c->d_Dsum, nb, cudaMemcpyDeviceToHost, c->stream));
  uint64_t n = 0;
  if (int r = read_accepts(c, &n)) return r;
  if (n_accepted) *n_accepted = n;
  EPV_CUDA(c, cudaGetLastError());
  return 0;
}

int epv_run_mcmc(epv_ctx *c, uint64_t burn_in, uint64_t batch, uint64_t seed,
                 uint64_t first_sweep, double *J, double *D, double *acc_rate) {
  uint64_t n = 0;
  if (int r = epv_run_mcmc_raw(c, burn_in, batch, seed, first_sweep, J, D, &n)) return r;
  for (int i = 0; i < c->n_nodes*8; ++i) {                   // batch average, :590-594
    J[i] /= batch; D[i] /= batch;
  }
  if (acc_rate) *acc_rate = (double)n/(batch*(c->n_global - 2));   // :597
  return 0;
}

int epv_suffstats(epv_ctx *c, double *J, double *D) {
  if (!c) return 1;
  if (!c->n_sites) return fail(c, "no paths loaded");
  cudaSetDevice(c->device);
  if (int r = do_suffstats(c)) return r;
  const size_t nb = (size_t)c->n_nodes*8*sizeof(double);
  EPV_CUDA(c, cudaMemcpyAsync(J, c->d_J, nb, cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaMemcpyAsync(D, c->d_D, nb, cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  EPV_CUDA(c, cudaGetLastError());
  return 0;
}

/* J/D statistics as exact integer accumulators, for drivers of several shards: epv_stats_begin clears
 * them, epv_stats_accumulate adds one J/D pass over the current paths (enqueued only), epv_stats_export
 * copies them (epv_stats_words 8-byte words) to a DEVICE buffer of the caller, which adds the shards'
 * buffers element-wise as 64-bit integers (peer copies + epv_stats_add, or an NCCL all-reduce on int64),
 * and epv_stats_fold turns a summed buffer (NULL: the context's own) into J / D [n_nodes*8].  Integer
 * sums are exact, so J and D do not depend on how the sites were sharded. */
int epv_stats_words(epv_ctx *c, uint64_t *words) {
  if (!c || !words) return 1;
  if (!c->n_nodes) return fail(c, "epv_set_tree must be called first");
  *words = suffstats_scratch_bytes(c->n_nodes, c->sm_count)/8;
  return 0;
}

int epv_stats_begin(epv_ctx *c) {
  if (!c) return 1;
  if (!c->n_sites) return fail(c, "no paths loaded");
  cudaSetDevice(c->device);
  EPV_CUDA(c, cudaMemsetAsync(c->d_ss, 0, suffstats_scratch_bytes(c->n_nodes, c->sm_count), c->stream));
  c->stats_valid = false;
  return 0;
}

int epv_stats_accumulate(epv_ctx *c) {
  if (!c) return 1;
  if (!c->n_sites) return fail(c, "no paths loaded");
  cudaSetDevice(c->device);
  return do_suffstats(c, false, false);
}

int epv_stats_export(epv_ctx *c, void *dev_words) {
  if (!c || !dev_words) return 1;
  cudaSetDevice(c->device);
  EPV_CUDA(c, cudaMemcpyAsync(dev_words, c->d_ss, suffstats_scratch_bytes(c->n_nodes, c->sm_count),
                              cudaMemcpyDeviceToDevice, c->stream));
  return 0;
}

int epv_stats_add(epv_ctx *c, const void *dev_words) {
  if (!c || !dev_words) return 1;
  cudaSetDevice(c->device);
  launch_suffstats_add(c->d_ss, static_cast<const unsigned long long *>(dev_words), c->n_nodes, c->stream);
  ++c->launches;
  return 0;
}

int epv_stats_fold(epv_ctx *c, const void *dev_words, double *J, double *D) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  const unsigned long long *src = dev_words ? static_cast<const unsigned long long *>(dev_words) : c->d_ss;
  launch_suffstats_fold(src, c->tree(), c->d_Jsum, c->d_Dsum, c->stream);
  ++c->launches;
  const size_t nb = (size_t)c->n_nodes*8*sizeof(double);
  EPV_CUDA(c, cudaMemcpyAsync(J, c->d_Jsum, nb, cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaMemcpyAsync(D, c->d_Dsum, nb, cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  EPV_CUDA(c, cudaGetLastError());
  return c->n_sites ? housekeeping_take(c) : 0;
}

/* The J/D pass enqueued only (a multi-GPU driver starts it on every shard before reading any back). */
int epv_suffstats_async(epv_ctx *c) {
  if (!c) return 1;
  if (!c->n_sites) return fail(c, "no paths loaded");
  cudaSetDevice(c->device);
  return do_suffstats(c);
}

int epv_suffstats_fetch(epv_ctx *c, double *J, double *D) {
  if (!c) return 1;
  if (!c->stats_valid) return fail(c, "epv_suffstats_fetch needs a preceding epv_suffstats_async on the current paths");
  cudaSetDevice(c->device);
  const size_t nb = (size_t)c->n_nodes*8*sizeof(double);
  EPV_CUDA(c, cudaMemcpyAsync(J, c->d_J, nb, cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaMemcpyAsync(D, c->d_D, nb, cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  EPV_CUDA(c, cudaGetLastError());
  return 0;
}

int epv_llh_grad(epv_ctx *c, const double *rates, double *llh, double *grad) {
  if (!c) return 1;
  if (!c->stats_valid) return fail(c, "epv_llh_grad needs a preceding epv_suffstats on the current paths");
  cudaSetDevice(c->device);
  launch_llh_grad(c->d_J, c->d_D, c->n_nodes, rates, c->d_out9, c->stream);
  ++c->launches;
  double out[9];
  EPV_CUDA(c, cudaMemcpyAsync(c->h_mail, c->d_out9, sizeof out, cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  std::memcpy(out, c->h_mail, sizeof out);
  *llh = out[0];
  for (int i = 0; i < 8; ++i) grad[i] = out[1 + i];
  return 0;
}

int epv_scale_jump_times(epv_ctx *c, const double *new_branches) {
  if (!c) return 1;
  if (!c->n_sites) return fail(c, "no paths loaded");
  cudaSetDevice(c->device);
  std::vector<double> scale(c->n_nodes, 1.0);
  for (int b = 1; b < c->n_nodes; ++b) {
    if (!(new_branches[b] > 0.0)) return fail(c, "branch lengths must be positive");
    scale[b] = new_branches[b]/c->h_branch[b];                // ParamEstimation.cpp:373
  }
  double *d_scale = reinterpret_cast<double *>(c->d_mail);
  EPV_CUDA(c, cudaMemcpyAsync(d_scale, scale.data(), c->n_nodes*sizeof(double), cudaMemcpyHostToDevice, c->stream));
  launch_scale_jumps(c->d_jmp[c->cur], c->d_jbase, c->d_off[c->cur], c->off_pitch, c->n_own(),
                     c->n_nodes - 1, d_scale, c->d_hjmp, c->halo_k, c->stream);
  c->launches += 2;
  for (int b = 1; b < c->n_nodes; ++b) c->h_branch[b] = new_branches[b];
  EPV_CUDA(c, cudaMemcpyAsync(c->d_branch, c->h_branch.data(), c->n_nodes*sizeof(double), cudaMemcpyHostToDevice, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  c->tri_valid = false; c->stats_valid = false;
  EPV_CUDA(c, cudaGetLastError());
  return housekeeping_take(c);
}

int epv_mcem_iteration(epv_ctx *c, uint64_t burn_in, uint64_t batch, uint64_t seed, uint64_t first_sweep,
                       int optimize_branches, double param_tol, double *rates, double *T,
                       double *branches, double *baseline, double *acc_rate, double *llh) {
  if (!c) return 1;
  if (!c->n_sites) return fail(c, "no paths loaded");
  for (int b = 1; b < c->n_nodes; ++b)
    if (branches[b] != c->h_branch[b])
      return fail(c, "epv_mcem_iteration: branches differ from the context's tree (paths' tot_time)");
  if (int r = epv_set_model(c, rates, T)) return r;          // reset follows from the new model
  std::vector<double> J((size_t)c->n_nodes*8), D((size_t)c->n_nodes*8);
  if (int r = epv_run_mcmc(c, burn_in, batch, seed, first_sweep, J.data(), D.data(), acc_rate)) return r;
  double bl[4];
  if (!optimize_branches) {                                  // epievo_est_params_histories.cpp:254-257
    *llh = epv_estimate_rates((uint64_t)c->n_nodes, J.data(), D.data(), param_tol, rates, T, bl);
    epv_set_one_change(rates, (uint64_t)c->n_nodes, branches);
  }
  else                                                       // :259-262
    *llh = epv_estimate_rates_and_branches((uint64_t)c->n_nodes, J.data(), D.data(), param_tol, rates,
                                           branches, T, bl);
  if (baseline) for (int i = 0; i < 4; ++i) baseline[i] = bl[i];
  return epv_scale_jump_times(c, branches);                  // :264
}

uint64_t epv_launch_count(const epv_ctx *c) { return c ? c->launches : 0; }
uint64_t epv_recovery_count(const epv_ctx *c) { return c ? c->recoveries : 0; }

int epv_enable_kernel_timing(epv_ctx *c, int on) {
  if (!c) return 1;
  if (!on) drain_timing(c);
  c->timing = on != 0;
  return 0;
}

int epv_rebuild_kernel_time(epv_ctx *c, double *ms) {
  if (!c || !ms) return 1;
  *ms = c->last_rebuild_ms;
  return 0;
}

int epv_mh_kernel_time(epv_ctx *c, int reset, double *ms, uint64_t *launches) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  drain_timing(c);
  if (ms) *ms = c->mh_ms;
  c->last_rebuild_ms = c->rebuild_ms;
  if (launches) *launches = c->mh_launches;
  if (reset) { c->mh_ms = 0.0; c->rebuild_ms = 0.0; c->mh_launches = 0; }
  return 0;
}

}  // extern "C"

extern "C" {

int epv_pass(epv_ctx *c, int colour, uint64_t seed, uint64_t sweep_idx) {
  if (int r = ready(c)) return r;
  if (colour < 0 || colour > 2) return fail(c, "colour must be 0, 1 or 2");
  if (!c->tri_valid) if (int r = do_reset(c)) return r;
  if (int r = one_pass(c, colour, seed, (uint32_t)sweep_idx)) return r;
  if (colour == 2) if (int r = housekeeping_lagged(c)) return r;
  EPV_CUDA(c, cudaGetLastError());
  return 0;
}

// The colour pass of a SHARDED context with the halo traffic overlapped (see one_pass): on return the
// side stream holds: edge update -> pack of the boundary sites' records into rec_lo / rec_hi; the main
// stream: interior update -> rebuild.  The driver moves the records on the side stream and calls
// epv_pass_end with what it received.
int epv_pass_begin(epv_ctx *c, int colour, uint64_t seed, uint64_t sweep_idx, void *rec_lo, void *rec_hi) {
  if (int r = ready(c)) return r;
  if (colour < 0 || colour > 2) return fail(c, "colour must be 0, 1 or 2");
  if (c->split_open) return fail(c, "epv_pass_begin: the previous pass was not closed with epv_pass_end");
  if (!c->tri_valid) if (int r = do_reset(c)) return r;
  if (int r = one_pass(c, colour, seed, (uint32_t)sweep_idx, nullptr, true)) return r;
  c->split_open = true;
  if (c->halo_lo || c->halo_hi) {
    const int64_t own_lo = (int64_t)c->halo_lo, own_hi = (int64_t)(c->n_sites - c->halo_hi);
    int64_t site_lo = -1, site_hi = -1;
    if (c->halo_lo) {
      if (!rec_lo) return fail(c, "epv_pass_begin: this shard has a low neighbour, rec_lo is required");
      site_lo = own_lo + colour;
    }
    if (c->halo_hi) {
      if (!rec_hi) return fail(c, "epv_pass_begin: this shard has a high neighbour, rec_hi is required");
      site_hi = own_hi - 3 + (((colour - (own_hi - 3)) % 3) + 3) % 3;
    }
    HaloArgs h;
    h.pv = c->view(c->cur ^ 1);                       // the CSR this pass read (one_pass has flipped cur)
    h.state = c->d_state; h.W = c->W; h.tri_llh = c->d_tri;
    h.n_sites = c->n_sites; h.global0 = c->global0();
    h.hcnt = c->d_hcnt; h.hjmp = c->d_hjmp; h.err = c->d_err; h.hmax = c->d_hmax;
    h.stage_off = c->d_stg_off; h.stage = c->d_stg;
    launch_halo_pack(h, site_lo, static_cast<unsigned long long *>(rec_lo), site_hi,
                     static_cast<unsigned long long *>(rec_hi), c->side_stream, true);
    ++c->launches;
  }
  EPV_CUDA(c, cudaGetLastError());
  return 0;
}

// Closes the pass: unpacks the neighbours' records (side stream, after whatever the driver enqueued
// there) and makes the main stream's next work wait for it.  After colour 2: housekeeping (one host
// synchronisation per sweep).
int epv_pass_end(epv_ctx *c, int colour, const void *rec_lo, const void *rec_hi) {
  if (!c) return 1;
  if (!c->split_open) return fail(c, "epv_pass_end without epv_pass_begin");
  cudaSetDevice(c->device);
  c->split_open = false;
  if (c->halo_lo || c->halo_hi) {
    const int64_t own_hi = (int64_t)(c->n_sites - c->halo_hi);
    int64_t site_lo = -1, site_hi = -1;
    if (c->halo_lo) {
      if (!rec_lo) return fail(c, "epv_pass_end: this shard has a low neighbour, rec_lo is required");
      site_lo = colour;
    }
    if (c->halo_hi) {
      if (!rec_hi) return fail(c, "epv_pass_end: this shard has a high neighbour, rec_hi is required");
      site_hi = own_hi + (((colour - own_hi) % 3) + 3) % 3;
    }
    HaloArgs h;
    h.pv = c->view(c->cur);
    h.state = c->d_state; h.W = c->W; h.tri_llh = c->d_tri;
    h.n_sites = c->n_sites; h.global0 = c->global0();
    h.hcnt = c->d_hcnt; h.hjmp = c->d_hjmp; h.err = c->d_err; h.hmax = c->d_hmax;
    h.stage_off = c->d_stg_off; h.stage = c->d_stg;
    launch_halo_unpack(h, site_lo, static_cast<const unsigned long long *>(rec_lo), site_hi,
                       static_cast<const unsigned long long *>(rec_hi), c->side_stream);
    ++c->launches;
    EPV_CUDA(c, cudaEventRecord(c->ev_halo, c->side_stream));
    EPV_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_halo, 0));
  }
  if (colour == 2) if (int r = housekeeping_lagged(c)) return r;
  EPV_CUDA(c, cudaGetLastError());
  return 0;
}

int epv_accept_count(epv_ctx *c, int reset, uint64_t *n_accepted) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  uint64_t n = 0;
  if (int r = read_accepts(c, &n)) return r;
  if (n_accepted) *n_accepted = n;
  if (reset) EPV_CUDA(c, cudaMemsetAsync(c->d_nacc, 0, sizeof(unsigned long long), c->stream));
  return 0;
}

}  // extern "C"

// ---- site-independent E-step of epievo_initialization (IndepSite.cpp) -------------------
namespace {

int indep_ready(epv_ctx *c, const double *rates) {
  if (!c) return 1;
  if (!c->n_sites) return fail(c, "no paths loaded");
  if (rates && !(rates[0] > 0.0 && rates[1] > 0.0)) return fail(c, "the two rates must be positive");
  cudaSetDevice(c->device);
  if (c->itab_nodes != c->n_nodes) {
    cudaFree(c->d_itab); cudaFree(c->d_iacc); c->d_itab = nullptr; c->d_iacc = nullptr;
    EPV_CUDA(c, cudaMalloc(&c->d_itab, (size_t)c->n_nodes*sizeof(IndepTab)));
    EPV_CUDA(c, cudaMalloc(&c->d_iacc, (size_t)4*c->n_nodes*sizeof(double)));
    c->itab_nodes = c->n_nodes;
  }
  return 0;
}

int indep_tables(epv_ctx *c, const double *rates) {
  std::vector<IndepTab> tab(c->n_nodes);
  std::memset(tab.data(), 0, tab.size()*sizeof(IndepTab));
  for (int v = 1; v < c->n_nodes; ++v)
    epv_indep_tables(rates[0], rates[1], c->h_branch[v], tab[v].P, tab[v].J0, tab[v].J1, tab[v].D0, tab[v].D1);
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));           // the previous table may still be in use
  EPV_CUDA(c, cudaMemcpyAsync(c->d_itab, tab.data(), tab.size()*sizeof(IndepTab), cudaMemcpyHostToDevice, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  return 0;
}

int indep_read(epv_ctx *c, double *J, double *D) {
  std::vector<double> acc((size_t)4*c->n_nodes);
  EPV_CUDA(c, cudaMemcpyAsync(acc.data(), c->d_iacc, acc.size()*sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  EPV_CUDA(c, cudaGetLastError());
  for (int v = 0; v < c->n_nodes; ++v) {
    J[2*v] = acc[4*v]; J[2*v + 1] = acc[4*v + 1];
    D[2*v] = acc[4*v + 2]; D[2*v + 1] = acc[4*v + 3];
  }
  return 0;
}

}  // namespace

extern "C" {

int epv_indep_suffstats(epv_ctx *c, double *J, double *D) {
  if (int r = indep_ready(c, nullptr)) return r;
  launch_indep_stats(c->view(c->cur), c->d_state, c->W, c->tree(), c->d_iacc, c->sm_count, c->stream);
  ++c->launches;
  return indep_read(c, J, D);
}

int epv_indep_expectation(epv_ctx *c, const double *rates, double *J, double *D) {
  if (int r = indep_ready(c, rates)) return r;
  if (int r = indep_tables(c, rates)) return r;
  const double pi0 = rates[1]/(rates[0] + rates[1]);       // root_post_prob0, IndepSite.cpp:94
  launch_indep_expect(c->view(c->cur), c->d_state, c->W, c->tree(), c->d_itab, pi0, c->d_iacc, c->sm_count,
                      c->stream);
  ++c->launches;
  return indep_read(c, J, D);
}

int epv_indep_update_paths(epv_ctx *c, const double *rates, uint64_t seed, uint64_t iteration) {
  if (int r = indep_ready(c, rates)) return r;
  if (int r = indep_tables(c, rates)) return r;
  IndepLaunch il; il.tab = c->d_itab; il.r0 = rates[0]; il.r1 = rates[1];
  for (int colour = 0; colour < 3; ++colour)
    if (int r = one_pass(c, colour, seed, (uint32_t)iteration, &il)) return r;
  c->tri_valid = false;
  if (int r = housekeeping(c)) return r;
  EPV_CUDA(c, cudaGetLastError());
  return 0;
}

}  // extern "C"
