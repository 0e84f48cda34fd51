// api_paths.cu -- host exchange format <-> device CSR, capacity management.
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "../../include/epievo_b200.h"
#include "context.h"

using namespace epv;

namespace epv {

// Current jump count of every branch, read from the tail of the offset arrays.
int read_branch_totals(epv_ctx *c, std::vector<uint64_t> &tot) {
  const int n_br = c->n_nodes - 1;
  tot.assign(n_br, 0);
  // one strided copy: the closing offset of every branch
  EPV_CUDA(c, cudaMemcpy2DAsync(c->h_mail, sizeof(uint32_t), c->d_off[c->cur] + c->n_own(), c->off_pitch*sizeof(uint32_t),
                                sizeof(uint32_t), (size_t)n_br, cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  uint64_t sum = 0;
  for (int bi = 0; bi < n_br; ++bi) sum += tot[bi] = reinterpret_cast<uint32_t *>(c->h_mail)[bi];
  c->jumps_per_row = (double)sum/((double)n_br*(double)c->n_own());
  return 0;
}

static uint64_t capacity_for(uint64_t count, uint64_t n_sites) {
  // room for the stationary fluctuation of the total plus growth between checks
  uint64_t cap = count + count/2 + n_sites/16 + 4096;
  // test-only: almost no headroom at upload, so that growth must go through the recovery of one_pass
  static const bool tight = std::getenv("EPIEVO_B200_TIGHT_CAPACITY") != nullptr;
  if (tight) cap = count/2 + count/32 + 64;            // (callers pass 2 x the fill)
  return (cap + 15) & ~15ull;
}

// Allocate both jump buffers for the given per-branch capacities.
static int alloc_jumps(epv_ctx *c, const std::vector<uint64_t> &cap, double *out[2],
                       std::vector<uint64_t> &base, uint64_t &total) {
  base.assign(cap.size(), 0);
  total = 0;
  for (size_t bi = 0; bi < cap.size(); ++bi) { base[bi] = total; total += cap[bi]; }
  for (int k = 0; k < 2; ++k) {
    out[k] = nullptr;
    EPV_CUDA(c, cudaMalloc(&out[k], std::max<uint64_t>(total, 1)*sizeof(double)));
  }
  return 0;
}

// Staging blob for the accepted proposals of ONE colour pass: at most one header per
// group (ceil(n_br/4) words of u16 counts) plus the new jump times of a third of the rows
// (sized at half of the total jump capacity).
// Global scratch for proposals longer than the update kernel's per-thread buffer (16 / 32 / 64 / 128 jump
// times by tree size): one slot per overflowing site of a pass.  Sized from the data: when a site's whole-tree
// proposal is expected to exceed half of that buffer, most sites overflow -> a slot per group, of 4 x the
// expectation + 64; otherwise a small pool of long slots.
static int ensure_scratch(epv_ctx *c, double t_sum, double r_mean) {
  const int maxj = c->n_nodes <= 8 ? 16 : c->n_nodes <= 16 ? 32 : c->n_nodes <= 64 ? 64 : 128;
  const double expect = std::max(c->jumps_per_row*(double)(c->n_nodes - 1), t_sum*r_mean);
  uint64_t slots, cap;
  if (2.0*expect > (double)maxj) { slots = c->n_groups + 64; cap = ((uint64_t)(4.0*expect) + 64 + 15) & ~15ull; if (cap < 256) cap = 256; }
  else { slots = std::max<uint64_t>(2048, c->n_groups/64); cap = 2048; }
  if (cap > 0x7fffffffull || slots > 0xfffffff0ull) return fail(c, "proposal scratch too large; use more shards");
  if (slots <= c->scratch_slots && cap <= c->scratch_cap) return 0;
  slots = std::max<uint64_t>(slots, c->scratch_slots); cap = std::max<uint64_t>(cap, c->scratch_cap);
  if (slots > c->n_groups + 64 && cap > 256) {       // the small pool of long slots became a slot per group: shorter slots
    cap = ((uint64_t)(4.0*expect) + 64 + 15) & ~15ull; if (cap < 256) cap = 256;
  }
  cudaFree(c->d_scratch); c->d_scratch = nullptr;
  EPV_CUDA(c, cudaMalloc(&c->d_scratch, slots*cap*sizeof(double)));
  c->scratch_slots = (uint32_t)slots; c->scratch_cap = (uint32_t)cap;
  return 0;
}

int alloc_stage(epv_ctx *c) {
  const uint64_t n_br = c->n_nodes - 1;
  c->n_groups = (c->n_sites + 2)/3;
  // jump times of one pass: half of the stored capacity, or twice what a third of the sites holds at
  // stationarity (sum of the branch lengths x the mean triplet rate per site) -- histories that start with
  // few jumps (epievo_initialization's heuristic paths) grow towards that within a sweep
  double t_sum = 0.0, r_mean = 1.0;
  for (int v = 1; v < c->n_nodes; ++v) t_sum += c->h_branch[v];
  if (c->have_model) { r_mean = 0.0; for (int i = 0; i < 8; ++i) r_mean += c->mp.rate[i]/8.0; if (r_mean < 1.0) r_mean = 1.0; }
  if (int r = ensure_scratch(c, t_sum, c->have_model ? 1.0 : 0.0)) return r;
  const unsigned long long model_jumps = (unsigned long long)(2.0*r_mean*t_sum*(double)c->n_groups);
  const unsigned long long cap = c->n_groups*((n_br + 3)/4) + std::max<unsigned long long>(c->jmp_total_cap/2, model_jumps) + 4096;
  if (cap >= 0xfffffff0ull)
    return fail(c, "the proposal staging of one pass would exceed 2^32 words on one GPU; use more shards");
  if (c->d_stg && cap <= c->stg_cap) return 0;
  cudaFree(c->d_stg); c->d_stg = nullptr;
  EPV_CUDA(c, cudaMalloc(&c->d_stg, cap*sizeof(double)));
  c->stg_cap = cap;
  if (!c->d_stg_off) EPV_CUDA(c, cudaMalloc(&c->d_stg_off, c->n_groups*sizeof(uint32_t)));
  if (!c->d_ovf_list) EPV_CUDA(c, cudaMalloc(&c->d_ovf_list, (c->n_groups + 1)*sizeof(uint32_t)));
  return 0;
}

// Thin headroom = a pass could overflow a branch: more than half full, or less room than 1.5 x what the
// branch holds at stationarity (about one change per site per unit time).  one_pass then checks the error
// flag after every pass and recovers (regrow + rerun of the rebuild) instead of once per sweep.
void update_guard(epv_ctx *c, const std::vector<uint64_t> &tot) {
  bool g = false;
  for (size_t bi = 0; bi < tot.size(); ++bi) {
    const double expect = 1.5*(double)c->n_own()*c->h_branch[bi + 1];
    if (2*tot[bi] > c->h_jcap[bi] || expect > (double)c->h_jcap[bi]) g = true;
  }
  c->guard = g;
}

// New jump buffers with the given capacities (>= the current ones); the CURRENT buffer's contents (tot[b]
// jumps per branch) are carried over.  keep_stage: leave the proposal staging alone (it is in use).
int regrow_to(epv_ctx *c, const std::vector<uint64_t> &cap, const std::vector<uint64_t> &tot, bool keep_stage) {
  const int n_br = c->n_nodes - 1;
  std::vector<uint64_t> base;
  double *nj[2]; uint64_t total;
  if (int r = alloc_jumps(c, cap, nj, base, total)) return r;
  for (int bi = 0; bi < n_br; ++bi)
    EPV_CUDA(c, cudaMemcpyAsync(nj[c->cur] + base[bi], c->d_jmp[c->cur] + c->h_jbase[bi],
                                tot[bi]*sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
  EPV_CUDA(c, cudaMemcpyAsync(c->d_jbase, base.data(), n_br*sizeof(uint64_t), cudaMemcpyHostToDevice, c->stream));
  EPV_CUDA(c, cudaMemcpyAsync(c->d_jcap, cap.data(), n_br*sizeof(uint64_t), cudaMemcpyHostToDevice, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  cudaFree(c->d_jmp[0]); cudaFree(c->d_jmp[1]);
  c->d_jmp[0] = nj[0]; c->d_jmp[1] = nj[1];
  c->h_jbase = base; c->h_jcap = cap; c->jmp_total_cap = total;
  update_guard(c, tot);
  return keep_stage ? 0 : alloc_stage(c);
}

// Grow any branch that is more than half full to three times its fill (called between sweeps).
int maybe_regrow(epv_ctx *c, const std::vector<uint64_t> &tot) {
  const int n_br = c->n_nodes - 1;
  bool need = false;
  for (int bi = 0; bi < n_br; ++bi)
    if (tot[bi] > 0xfffffff0ull) return fail(c, "a branch holds more than 2^32 jumps on one GPU; use more shards");
    else if (2*tot[bi] > c->h_jcap[bi]) need = true;
  if (!need) { update_guard(c, tot); return alloc_stage(c); }
  std::vector<uint64_t> cap(n_br);
  for (int bi = 0; bi < n_br; ++bi) cap[bi] = std::max(c->h_jcap[bi], capacity_for(2*tot[bi], c->n_own()));
  return regrow_to(c, cap, tot, false);
}


}  // namespace epv

namespace epv {

// Host copy of the halo store: counts [2*HALO_W][n_br] and jumps [..][halo_k].
int read_halo_store(epv_ctx *c, std::vector<uint16_t> &cnt, std::vector<double> &jmp) {
  const size_t n_br = c->n_nodes - 1, rows = 2*HALO_W*n_br;
  cnt.assign(rows, 0); jmp.assign(rows*c->halo_k, 0.0);
  EPV_CUDA(c, cudaMemcpyAsync(cnt.data(), c->d_hcnt, rows*sizeof(uint16_t), cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaMemcpyAsync(jmp.data(), c->d_hjmp, rows*c->halo_k*sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  return 0;
}

// Row capacity of the halo store for boundary rows of up to kmax jumps now: twice that, and four times what a
// row of the longest branch holds at stationarity (about 1.5 changes per unit time, as update_guard) -- histories
// that start short on long branches (T = 20) grow several-fold in their first sweeps, before any regrow rule,
// read late or not, has seen them.
int halo_capacity_for(const epv_ctx *c, uint64_t kmax) {
  double t_max = 0.0;
  for (int b = 1; b < c->n_nodes; ++b) t_max = std::max(t_max, c->h_branch[b]);
  // test-only: no headroom over the rows present, so that growth has to come from the between-sweeps check
  if (std::getenv("EPIEVO_B200_TIGHT_HALO")) return (int)std::max<uint64_t>(HALO_K_MIN, kmax + 8);
  const double want = std::max({(double)HALO_K_MIN, 2.0*(double)kmax + 16.0, 6.0*t_max + 16.0});
  return (int)std::min(want, 60000.0);
}

// Halo rows hold up to k jumps each from now on (k only grows); the store's contents are kept.
int set_halo_capacity(epv_ctx *c, int k) {
  k = (k + 7) & ~7;
  if (k <= c->halo_k && c->d_hjmp) return 0;
  if (k < c->halo_k) k = c->halo_k;
  const size_t n_br = c->n_nodes - 1, rows = 2*HALO_W*n_br;
  std::vector<uint16_t> cnt; std::vector<double> old;
  const int k_old = c->halo_k;
  if (c->d_hjmp) {
    if (c->side_stream) EPV_CUDA(c, cudaStreamSynchronize(c->side_stream));
    if (int r = read_halo_store(c, cnt, old)) return r;
    cudaFree(c->d_hjmp); c->d_hjmp = nullptr;
  }
  EPV_CUDA(c, cudaMalloc(&c->d_hjmp, rows*k*sizeof(double)));
  c->halo_k = k;
  if (!old.empty()) {
    std::vector<double> nw(rows*k, 0.0);
    for (size_t r = 0; r < rows; ++r) std::memcpy(&nw[r*k], &old[r*k_old], (size_t)cnt[r]*sizeof(double));
    EPV_CUDA(c, cudaMemcpy(c->d_hjmp, nw.data(), nw.size()*sizeof(double), cudaMemcpyHostToDevice));
  }
  return 0;
}

// Local site of halo slot k (0..2 low halo, 3..5 high halo), or -1 if the slot is unused.
int64_t slot_site(const epv_ctx *c, int k) {
  if (k < HALO_W) return (uint64_t)k < c->halo_lo ? k : -1;
  return (uint64_t)(k - HALO_W) < c->halo_hi ? (int64_t)(c->n_sites - c->halo_hi) + (k - HALO_W) : -1;
}

}  // namespace epv

namespace epv {

// Argument checks shared by the two upload entry points; fills in the default shard.
int check_upload(epv_ctx *c, uint64_t n_sites, uint64_t n_nodes, const double *tot_time) {
  if (!c->n_nodes) return fail(c, "epv_set_tree must be called before epv_upload_paths");
  if ((int)n_nodes != c->n_nodes) return fail(c, "paths and tree disagree on the number of nodes");
  if (n_sites < 3) return fail(c, "need at least 3 sites");
  cudaSetDevice(c->device);
  if (!c->have_shard) {
    c->n_global = n_sites; c->site_begin = 0; c->site_end = n_sites; c->halo_lo = c->halo_hi = 0;
  }
  if (n_sites != c->site_end - c->site_begin + c->halo_lo + c->halo_hi)
    return fail(c, "n_sites does not match the shard set with epv_set_shard");
  for (int b = 1; b < c->n_nodes; ++b)
    if (tot_time[b] != c->h_branch[b])
      return fail(c, "paths' tot_time differs from the tree's branch length (the reference "
                     "asserts equality, Segment.cpp:76); rescale the paths or the tree");
  // new paths: a device-side failure of the previous ones (sticky flags) no longer applies
  EPV_CUDA(c, cudaMemsetAsync(c->d_err, 0, sizeof(unsigned), c->stream));
  EPV_CUDA(c, cudaMemsetAsync(c->d_hmax, 0, sizeof(unsigned), c->stream));
  c->hk_pending = false;                 // a late housekeeping record of the previous paths means nothing now
  return 0;
}

// Everything whose size depends only on (n_sites, n_nodes); kept across uploads of the
// same shape.
int ensure_shape(epv_ctx *c, uint64_t n_sites) {
  if (c->n_sites == n_sites && c->d_off[0]) return 0;
  free_paths(c);
  c->h_jcap.clear();
  const int n_br = c->n_nodes - 1;
  const uint64_t n_own = n_sites - c->halo_lo - c->halo_hi;
  c->n_sites = n_sites;
  c->W = 1;                                         // bytes of state bits per site, a power of two
  while (c->W*8 < c->n_nodes) c->W *= 2;            // so that <= 64 nodes are one aligned load
  c->off_pitch = (n_own + 1 + 4 + 31) & ~31ull;     // room for the rebuild kernel's 16-byte vectors
  c->n_tiles = (uint32_t)((n_own + 1 + RB_TILE_ROWS - 1)/RB_TILE_ROWS);
  for (int k = 0; k < 2; ++k)
    EPV_CUDA(c, cudaMalloc(&c->d_off[k], (size_t)n_br*c->off_pitch*sizeof(uint32_t)));
  EPV_CUDA(c, cudaMalloc(&c->d_jbase, n_br*sizeof(uint64_t)));
  EPV_CUDA(c, cudaMalloc(&c->d_jcap, n_br*sizeof(uint64_t)));
  EPV_CUDA(c, cudaMalloc(&c->d_jneed, n_br*sizeof(unsigned long long)));
  EPV_CUDA(c, cudaMemsetAsync(c->d_jneed, 0, n_br*sizeof(unsigned long long), c->stream));
  EPV_CUDA(c, cudaMalloc(&c->d_state, (size_t)n_sites*c->W));
  EPV_CUDA(c, cudaMalloc(&c->d_tri, n_sites*sizeof(double)));
  EPV_CUDA(c, cudaMalloc(&c->d_desc, (size_t)n_br*c->n_tiles*sizeof(unsigned long long)));
  EPV_CUDA(c, cudaMalloc(&c->d_hcnt, (size_t)2*HALO_W*n_br*sizeof(uint16_t)));
  if (int r = set_halo_capacity(c, c->halo_k)) return r;
  const uint64_t n_rows = (uint64_t)c->n_nodes*n_sites;
  EPV_CUDA(c, cudaMalloc(&c->d_stage_init, n_rows));
  EPV_CUDA(c, cudaMalloc(&c->d_stage_off, (n_rows + 1)*sizeof(unsigned long long)));
  EPV_CUDA(c, cudaMemsetAsync(c->d_desc, 0, (size_t)n_br*c->n_tiles*sizeof(unsigned long long), c->stream));
  c->epoch = 0;
  return 0;
}

// Jump buffers (+ the proposal staging blob): reallocated only when a branch no longer fits.
int ensure_jumps(epv_ctx *c, const std::vector<uint64_t> &count) {
  const int n_br = c->n_nodes - 1;
  uint64_t sum = 0;
  for (int bi = 0; bi < n_br; ++bi) sum += count[bi];
  c->jumps_per_row = (double)sum/((double)n_br*(double)c->n_own());
  bool fits = (int)c->h_jcap.size() == n_br && c->d_jmp[0];
  for (int bi = 0; fits && bi < n_br; ++bi) fits = 2*count[bi] <= c->h_jcap[bi];
  for (int bi = 0; bi < n_br; ++bi)
    if (count[bi] > 0xfffffff0ull) return fail(c, "a branch holds more than 2^32 jumps on one GPU; use more shards");
  if (fits) { update_guard(c, count); return alloc_stage(c); }
  for (int k = 0; k < 2; ++k) { cudaFree(c->d_jmp[k]); c->d_jmp[k] = nullptr; }
  std::vector<uint64_t> cap(n_br);
  for (int bi = 0; bi < n_br; ++bi) cap[bi] = capacity_for(2*count[bi], c->n_own());
  if (int r = alloc_jumps(c, cap, c->d_jmp, c->h_jbase, c->jmp_total_cap)) return r;
  c->h_jcap = cap;
  update_guard(c, count);
  if (int r = alloc_stage(c)) return r;
  EPV_CUDA(c, cudaMemcpyAsync(c->d_jbase, c->h_jbase.data(), n_br*sizeof(uint64_t), cudaMemcpyHostToDevice, c->stream));
  EPV_CUDA(c, cudaMemcpyAsync(c->d_jcap, c->h_jcap.data(), n_br*sizeof(uint64_t), cudaMemcpyHostToDevice, c->stream));
  return 0;
}

}  // namespace epv

extern "C" {

int epv_upload_paths(epv_ctx *c, uint64_t n_sites, uint64_t n_nodes, const uint8_t *init_state,
                     const uint64_t *row_off, const double *jumps, const double *tot_time) {
  if (!c) return 1;
  if (int r = check_upload(c, n_sites, n_nodes, tot_time)) return r;
  const int n_br = c->n_nodes - 1;
  if (row_off[0] != 0 || row_off[n_sites] != row_off[0]) return fail(c, "root rows must be empty");
  const uint64_t own_lo = c->halo_lo, own_hi = n_sites - c->halo_hi, n_own = own_hi - own_lo;
  std::vector<uint64_t> count(n_br), first(n_br);
  for (int bi = 0; bi < n_br; ++bi) {
    const uint64_t r = (uint64_t)(bi + 1)*n_sites;
    if (row_off[r + n_sites] < row_off[r]) return fail(c, "row_off must be non-decreasing");
    first[bi] = row_off[r + own_lo];
    count[bi] = row_off[r + own_hi] - first[bi];
  }
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  if (int r = ensure_shape(c, n_sites)) return r;
  if (int r = ensure_jumps(c, count)) return r;
  c->cur = 0; c->tri_valid = false; c->stats_valid = false;

  // raw host arrays -> device staging -> layout conversion on the device
  const uint64_t n_rows = n_nodes*n_sites;
  EPV_CUDA(c, cudaMemcpyAsync(c->d_stage_init, init_state, n_rows, cudaMemcpyHostToDevice, c->stream));
  EPV_CUDA(c, cudaMemcpyAsync(c->d_stage_off, row_off, (n_rows + 1)*sizeof(unsigned long long), cudaMemcpyHostToDevice, c->stream));
  launch_import(c->d_stage_init, c->d_stage_off, n_sites, c->n_nodes, own_lo, n_own, c->d_off[0],
                c->off_pitch, c->d_state, c->W, c->stream);
  ++c->launches;
  for (int bi = 0; bi < n_br; ++bi)
    if (count[bi])
      EPV_CUDA(c, cudaMemcpyAsync(c->d_jmp[0] + c->h_jbase[bi], jumps + first[bi],
                                  count[bi]*sizeof(double), cudaMemcpyHostToDevice, c->stream));
  // halo rows -> halo store, its row capacity sized from the rows next to the shard's ends
  if (c->halo_lo || c->halo_hi) {
    uint64_t kmax = 0;
    for (int bi = 0; bi < n_br; ++bi)
      for (uint64_t d = 0; d < 2*HALO_W; ++d) {
        const uint64_t r = (uint64_t)(bi + 1)*n_sites;
        if (c->halo_lo) kmax = std::max(kmax, row_off[r + d + 1] - row_off[r + d]);
        if (c->halo_hi) kmax = std::max(kmax, row_off[r + n_sites - d] - row_off[r + n_sites - d - 1]);
      }
    if (kmax > 30000) return fail(c, "a shard-boundary row holds more than 30000 jumps");
    if (int r = set_halo_capacity(c, halo_capacity_for(c, kmax))) return r;
  }
  const size_t HK = c->halo_k;
  std::vector<uint16_t> hc((size_t)2*HALO_W*n_br, 0);
  std::vector<double> hj((size_t)2*HALO_W*n_br*HK, 0.0);
  for (int k = 0; k < 2*HALO_W; ++k) {
    const int64_t s = slot_site(c, k);
    if (s < 0) continue;
    for (int bi = 0; bi < n_br; ++bi) {
      const uint64_t r = (uint64_t)(bi + 1)*n_sites + (uint64_t)s;
      const uint64_t n = row_off[r + 1] - row_off[r];
      hc[(size_t)k*n_br + bi] = (uint16_t)n;
      std::memcpy(&hj[((size_t)k*n_br + bi)*HK], jumps + row_off[r], n*sizeof(double));
    }
  }
  EPV_CUDA(c, cudaMemcpyAsync(c->d_hcnt, hc.data(), hc.size()*sizeof(uint16_t), cudaMemcpyHostToDevice, c->stream));
  EPV_CUDA(c, cudaMemcpyAsync(c->d_hjmp, hj.data(), hj.size()*sizeof(double), cudaMemcpyHostToDevice, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  EPV_CUDA(c, cudaGetLastError());
  return 0;
}

// Jumps held per branch: halo-low rows, owned rows, halo-high rows.
static int branch_jump_layout(epv_ctx *c, std::vector<uint64_t> &own, std::vector<uint16_t> &hc,
                              std::vector<double> &hj, std::vector<uint64_t> &prefix) {
  const int n_br = c->n_nodes - 1;
  if (int r = read_branch_totals(c, own)) return r;
  if (int r = read_halo_store(c, hc, hj)) return r;
  prefix.assign(n_br + 1, 0);
  for (int bi = 0; bi < n_br; ++bi) {
    uint64_t t = own[bi];
    for (int k = 0; k < 2*HALO_W; ++k) if (slot_site(c, k) >= 0) t += hc[(size_t)k*n_br + bi];
    prefix[bi + 1] = prefix[bi] + t;
  }
  return 0;
}

// Jump times of every branch to the host array: low-halo rows, owned rows, high-halo rows.
static int copy_jumps_out(epv_ctx *c, const std::vector<uint64_t> &own, const std::vector<uint16_t> &hc,
                          const std::vector<double> &hj, const std::vector<uint64_t> &prefix,
                          double *jumps) {
  const int n_br = c->n_nodes - 1;
  for (int bi = 0; bi < n_br; ++bi) {
    double *dst = jumps + prefix[bi];
    for (int k = 0; k < HALO_W; ++k)
      if (slot_site(c, k) >= 0) {
        const size_t row = (size_t)k*n_br + bi;
        std::memcpy(dst, &hj[row*c->halo_k], hc[row]*sizeof(double)); dst += hc[row];
      }
    if (own[bi])
      EPV_CUDA(c, cudaMemcpyAsync(dst, c->d_jmp[c->cur] + c->h_jbase[bi], own[bi]*sizeof(double),
                                  cudaMemcpyDeviceToHost, c->stream));
    dst += own[bi];
    for (int k = HALO_W; k < 2*HALO_W; ++k)
      if (slot_site(c, k) >= 0) {
        const size_t row = (size_t)k*n_br + bi;
        std::memcpy(dst, &hj[row*c->halo_k], hc[row]*sizeof(double)); dst += hc[row];
      }
  }
  return 0;
}

int epv_paths_size(epv_ctx *c, uint64_t *n_sites, uint64_t *n_nodes, uint64_t *n_jumps) {
  if (!c) return 1;
  if (!c->n_sites) return fail(c, "no paths loaded");
  cudaSetDevice(c->device);
  std::vector<uint64_t> own, prefix; std::vector<uint16_t> hc; std::vector<double> hj;
  if (int r = branch_jump_layout(c, own, hc, hj, prefix)) return r;
  if (n_sites) *n_sites = c->n_sites;
  if (n_nodes) *n_nodes = (uint64_t)c->n_nodes;
  if (n_jumps) *n_jumps = prefix.back();
  return 0;
}

int epv_jump_counts(epv_ctx *c, uint64_t *per_node) {
  if (!c) return 1;
  if (!c->n_sites) return fail(c, "no paths loaded");
  cudaSetDevice(c->device);
  std::vector<uint64_t> tot;
  if (int r = read_branch_totals(c, tot)) return r;
  per_node[0] = 0;
  for (int bi = 0; bi + 1 < c->n_nodes; ++bi) per_node[bi + 1] = tot[bi];   // owned rows only
  return 0;
}

int epv_download_paths(epv_ctx *c, uint8_t *init_state, uint64_t *row_off, double *jumps,
                       double *tot_time) {
  if (!c) return 1;
  if (!c->n_sites) return fail(c, "no paths loaded");
  cudaSetDevice(c->device);
  if (int r = check_device_flags(c)) return r;
  const int n_br = c->n_nodes - 1;
  std::vector<uint64_t> own, prefix; std::vector<uint16_t> hc; std::vector<double> hj;
  if (int r = branch_jump_layout(c, own, hc, hj, prefix)) return r;
  const uint64_t n_rows = (uint64_t)c->n_nodes*c->n_sites;
  uint64_t *t_pre = reinterpret_cast<uint64_t *>(c->d_mail);
  if ((size_t)n_br + 1 > 4096) return fail(c, "too many branches");
  EPV_CUDA(c, cudaMemcpyAsync(t_pre, prefix.data(), (n_br + 1)*sizeof(uint64_t), cudaMemcpyHostToDevice, c->stream));
  launch_export(c->view(c->cur), c->d_state, c->W, c->n_sites, c->n_nodes, t_pre,
                c->d_stage_init, c->d_stage_off, c->stream);
  ++c->launches;
  EPV_CUDA(c, cudaMemcpyAsync(init_state, c->d_stage_init, n_rows, cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaMemcpyAsync(row_off, c->d_stage_off, (n_rows + 1)*sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
  if (int r = copy_jumps_out(c, own, hc, hj, prefix, jumps)) return r;
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  for (int b = 0; b < c->n_nodes; ++b) tot_time[b] = c->h_branch[b];
  EPV_CUDA(c, cudaGetLastError());
  return 0;
}

int epv_upload_paths_compact(epv_ctx *c, uint64_t n_sites, uint64_t n_nodes, const uint16_t *row_code,
                             const double *jumps, uint64_t n_jumps, const double *tot_time) {
  if (!c) return 1;
  if (int r = check_upload(c, n_sites, n_nodes, tot_time)) return r;
  const int n_br = c->n_nodes - 1;
  if ((size_t)3*n_br > 4096) return fail(c, "too many branches");
  const uint64_t own_lo = c->halo_lo, own_hi = n_sites - c->halo_hi, n_own = own_hi - own_lo;
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  if (int r = ensure_shape(c, n_sites)) return r;
  c->cur = 0; c->tri_valid = false; c->stats_valid = false;

  // codes -> device; per-branch totals come back so that the host can size the jump buffers
  const uint64_t n_rows = n_nodes*n_sites;
  uint16_t *d_code = reinterpret_cast<uint16_t *>(c->d_stage_off);
  EPV_CUDA(c, cudaMemcpyAsync(d_code, row_code, n_rows*sizeof(uint16_t), cudaMemcpyHostToDevice, c->stream));
  launch_compact_totals(d_code, n_sites, n_br, own_lo, own_hi, c->d_mail, c->stream);
  EPV_CUDA(c, cudaMemcpyAsync(c->h_mail, c->d_mail, (size_t)3*n_br*sizeof(unsigned long long),
                              cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  std::vector<uint64_t> count(n_br), first(n_br);
  uint64_t run = 0;
  for (int bi = 0; bi < n_br; ++bi) {
    first[bi] = run + c->h_mail[3*bi];
    count[bi] = c->h_mail[3*bi + 1];
    run = first[bi] + count[bi] + c->h_mail[3*bi + 2];
  }
  if (run != n_jumps) return fail(c, "the row codes' jump counts do not add up to n_jumps");
  if (int r = ensure_jumps(c, count)) return r;

  if (++c->epoch >= (1u << 30)) {
    EPV_CUDA(c, cudaMemsetAsync(c->d_desc, 0, (size_t)n_br*c->n_tiles*sizeof(unsigned long long), c->stream));
    c->epoch = 1;
  }
  launch_compact_import(d_code, n_sites, c->n_nodes, own_lo, n_own, c->d_off[0], c->off_pitch, c->d_state,
                        c->W, c->d_desc, c->n_tiles, c->epoch, c->d_ticket, c->ticket_base, c->stream);
  c->ticket_base += c->n_tiles;
  c->launches += 2;
  for (int bi = 0; bi < n_br; ++bi)
    if (count[bi])
      EPV_CUDA(c, cudaMemcpyAsync(c->d_jmp[0] + c->h_jbase[bi], jumps + first[bi],
                                  count[bi]*sizeof(double), cudaMemcpyHostToDevice, c->stream));
  // halo sites (at most 2*HALO_W): state bits and the fixed-capacity halo store, from the host
  if (c->halo_lo || c->halo_hi) {
    uint64_t kmax = 0;
    for (int bi = 0; bi < n_br; ++bi) {
      const uint16_t *row = row_code + (uint64_t)(bi + 1)*n_sites;
      for (uint64_t d = 0; d < 2*HALO_W; ++d) {
        if (c->halo_lo) kmax = std::max<uint64_t>(kmax, row[d] & 0x7fffu);
        if (c->halo_hi) kmax = std::max<uint64_t>(kmax, row[n_sites - 1 - d] & 0x7fffu);
      }
    }
    if (kmax > 30000) return fail(c, "a shard-boundary row holds more than 30000 jumps");
    if (int r = set_halo_capacity(c, halo_capacity_for(c, kmax))) return r;
  }
  const size_t HK = c->halo_k;
  std::vector<uint16_t> hc((size_t)2*HALO_W*n_br, 0);
  std::vector<double> hj((size_t)2*HALO_W*n_br*HK, 0.0);
  std::vector<uint8_t> hs((size_t)2*HALO_W*c->W, 0);
  for (int bi = -1; bi < n_br; ++bi) {
    const uint16_t *row = row_code + (uint64_t)(bi + 1)*n_sites;
    uint64_t lo = bi < 0 ? 0 : first[bi] - c->h_mail[3*bi], hi = bi < 0 ? 0 : first[bi] + count[bi];
    for (int k = 0; k < 2*HALO_W; ++k) {
      const int64_t s = slot_site(c, k);
      if (s < 0) continue;
      if (row[s] >> 15) hs[(size_t)k*c->W + (bi + 1)/8] |= (uint8_t)(1u << ((bi + 1) & 7));
      if (bi < 0) continue;
      const uint64_t n = row[s] & 0x7fffu;
      uint64_t &pos = k < HALO_W ? lo : hi;
      hc[(size_t)k*n_br + bi] = (uint16_t)n;
      std::memcpy(&hj[((size_t)k*n_br + bi)*HK], jumps + pos, n*sizeof(double));
      pos += n;
    }
  }
  for (int k = 0; k < 2*HALO_W; ++k) {
    const int64_t s = slot_site(c, k);
    if (s >= 0)
      EPV_CUDA(c, cudaMemcpyAsync(c->d_state + (uint64_t)s*c->W, &hs[(size_t)k*c->W], c->W,
                                  cudaMemcpyHostToDevice, c->stream));
  }
  EPV_CUDA(c, cudaMemcpyAsync(c->d_hcnt, hc.data(), hc.size()*sizeof(uint16_t), cudaMemcpyHostToDevice, c->stream));
  EPV_CUDA(c, cudaMemcpyAsync(c->d_hjmp, hj.data(), hj.size()*sizeof(double), cudaMemcpyHostToDevice, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  EPV_CUDA(c, cudaGetLastError());
  return 0;
}

int epv_download_paths_compact(epv_ctx *c, uint16_t *row_code, double *jumps, uint64_t jumps_cap,
                               uint64_t *n_jumps, double *tot_time) {
  if (!c) return 1;
  if (!c->n_sites) return fail(c, "no paths loaded");
  cudaSetDevice(c->device);
  if (int r = check_device_flags(c)) return r;
  const int n_br = c->n_nodes - 1;
  std::vector<uint64_t> own, prefix; std::vector<uint16_t> hc; std::vector<double> hj;
  if (int r = branch_jump_layout(c, own, hc, hj, prefix)) return r;
  if (n_jumps) *n_jumps = prefix.back();
  if (prefix.back() > jumps_cap) return fail(c, "jumps buffer too small (see *n_jumps / epv_paths_size)");
  const uint64_t n_rows = (uint64_t)c->n_nodes*c->n_sites;
  uint16_t *d_code = reinterpret_cast<uint16_t *>(c->d_stage_off);
  launch_compact_export(c->view(c->cur), c->d_state, c->W, c->n_sites, c->n_nodes, d_code, c->d_err, c->stream);
  ++c->launches;
  EPV_CUDA(c, cudaMemcpyAsync(row_code, d_code, n_rows*sizeof(uint16_t), cudaMemcpyDeviceToHost, c->stream));
  if (int r = copy_jumps_out(c, own, hc, hj, prefix, jumps)) return r;
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  for (int b = 0; b < c->n_nodes; ++b) tot_time[b] = c->h_branch[b];
  EPV_CUDA(c, cudaGetLastError());
  return check_device_flags(c);
}

}  // extern "C"
