// api_sim.cu -- forward simulation of histories on the device.
//
// Replaces epievo_sim's per-branch loop (prog/epievo_sim.cpp:102-152, :326-356 -- a global
// Gillespie simulation over the TripletSampler, one event of the whole sequence at a time)
// followed by global_jumps_to_paths.  Same law, no global event queue:
//
//   * thinning: every interior site carries candidate events from a Poisson process of rate
//     Lambda = max triplet rate; a candidate at (site s, time t) flips the site with probability
//     rate[x_{s-1}(t) x_s(t) x_{s+1}(t)] / Lambda.  The accepted flips are exactly the context-
//     dependent CTMC the reference simulates (end sites never change, TripletSampler.cpp:58-60);
//   * a candidate can be decided as soon as both neighbours have no undecided candidate before
//     it, i.e. when its time is a local minimum of the per-site "next candidate" times.  All
//     local minima are decided at once; a CTA iterates on its tile of sites in shared memory
//     until the tile is blocked by its two halo sites (whose values, frozen at launch, are lower
//     bounds), and the launch is repeated until no candidate is left.  The outcome does not
//     depend on the schedule: candidate times and acceptance draws come from Philox keyed by
//     (seed; candidate number, site, node);
//   * accepted jump times go to a fixed-capacity per-site buffer and are compacted into the
//     host exchange format (init_state [node][site], row_off, jumps) by a three-kernel scan.
//
// The result stays on the device until epv_simulate_fetch copies it out; feeding it to
// epv_upload_paths gives the sampler its starting histories.
#include <cmath>
#include <cstring>
#include <vector>

#include "../../include/epievo_b200.h"
#include "context.h"

namespace epv {

namespace {

constexpr int SIM_TILE = 512;                      // sites (= threads) per CTA
constexpr int SCAN_THREADS = 256, SCAN_ITEMS = 8;  // 2048 sites per CTA
constexpr unsigned long long PK_INF = 0x7ff0000000000000ull;
constexpr unsigned SIM_ERR_CAPACITY = 1u;

struct SimRates { double rate[8]; double lambda, inv_lambda; };

__device__ __forceinline__ double pk_time(unsigned long long pk) { return __longlong_as_double((long long)(pk & ~1ull)); }

// first candidate time of a site on a branch
__device__ __forceinline__ unsigned long long
first_candidate(uint64_t seed, uint64_t gsite, int node, int x, double inv_lambda, double T) {
  Philox rng; rng.init(seed, 0xffffffffu, gsite, (uint32_t)node);
  double u1, u2; rng.uniform2(u1, u2);
  const double tn = -log(u2)*inv_lambda;
  return tn < T ? (((unsigned long long)__double_as_longlong(tn) & ~1ull) | (unsigned)x) : (PK_INF | (unsigned)x);
}

// start of a branch: every site takes its parent's end state and its first candidate time
__global__ void
sim_init_kernel(uint64_t n, uint64_t site0, int node, const uint8_t *start, uint8_t *init_out,
                unsigned long long *pk, uint32_t *kc, uint32_t *cnt, uint64_t seed, double inv_lambda,
                double T) {
  const uint64_t s = (uint64_t)blockIdx.x*blockDim.x + threadIdx.x;
  if (s >= n) return;
  const int x = start[s];
  init_out[s] = (uint8_t)x;
  kc[s] = 0; cnt[s] = 0;
  pk[s] = (s == 0 || s + 1 == n) ? (PK_INF | (unsigned)x)
                                 : first_candidate(seed, site0 + s, node, x, inv_lambda, T);
}

__global__ void __launch_bounds__(SIM_TILE)
sim_round_kernel(uint64_t n, uint64_t site0, int node, unsigned long long *pk, uint32_t *kc, uint32_t *cnt,
                 double *jbuf, int K, const SimRates R, double T, uint64_t seed, unsigned *active,
                 unsigned *err) {
  __shared__ unsigned long long s_pk[SIM_TILE + 2];
  const int t = threadIdx.x;
  const uint64_t base = (uint64_t)blockIdx.x*SIM_TILE, s = base + t;
  unsigned long long my = s < n ? pk[s] : PK_INF;
  s_pk[t + 1] = my;
  if (t == 0) s_pk[0] = base > 0 ? pk[base - 1] : PK_INF;
  if (t == SIM_TILE - 1) s_pk[SIM_TILE + 1] = base + SIM_TILE < n ? pk[base + SIM_TILE] : PK_INF;
  uint32_t my_k = 0, my_cnt = 0;
  const bool live = s < n && my < PK_INF;
  if (live) { my_k = kc[s]; my_cnt = cnt[s]; }
  __syncthreads();
  for (;;) {
    const unsigned long long l = s_pk[t], r = s_pk[t + 2];
    // order of the candidates: by time, ties by site
    const bool res = my < PK_INF && pk_time(my) < pk_time(l) && pk_time(my) <= pk_time(r);
    if (res) {
      const double tc = pk_time(my);
      int x = (int)(my & 1);
      Philox rng; rng.init(seed, my_k, site0 + s, (uint32_t)node);
      double ua, ug; rng.uniform2(ua, ug);
      if (ua*R.lambda < R.rate[(int)(l & 1)*4 + x*2 + (int)(r & 1)]) {
        if ((int)my_cnt < K) jbuf[s*(uint64_t)K + my_cnt] = tc;
        else atomicOr(err, SIM_ERR_CAPACITY);
        ++my_cnt; x ^= 1;
      }
      ++my_k;
      const double tn = tc - log(ug)*R.inv_lambda;
      my = tn < T ? (((unsigned long long)__double_as_longlong(tn) & ~1ull) | (unsigned)x) : (PK_INF | (unsigned)x);
    }
    const int any = __syncthreads_or(res);       // all reads of this round are done
    if (!any) break;
    if (res) s_pk[t + 1] = my;
    __syncthreads();
  }
  if (live) {
    pk[s] = my; kc[s] = my_k; cnt[s] = my_cnt;
    if (my < PK_INF) *active = 1u;
  }
}

// end states of the branch (children start from them)
__global__ void
sim_end_kernel(uint64_t n, const unsigned long long *pk, uint8_t *end_out) {
  const uint64_t s = (uint64_t)blockIdx.x*blockDim.x + threadIdx.x;
  if (s < n) end_out[s] = (uint8_t)(pk[s] & 1);
}

// ---- counts -> row offsets, jump buffers -> one array: block sums, scan of the sums, scatter ----
__global__ void __launch_bounds__(SCAN_THREADS)
sim_block_sums_kernel(uint64_t n, const uint32_t *cnt, unsigned long long *partial) {
  __shared__ unsigned long long s_w[SCAN_THREADS/32];
  const uint64_t first = ((uint64_t)blockIdx.x*SCAN_THREADS + threadIdx.x)*SCAN_ITEMS;
  unsigned long long sum = 0;
  for (int k = 0; k < SCAN_ITEMS; ++k) if (first + k < n) sum += cnt[first + k];
  sum = warp_sum(sum);
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long tot = 0;
    for (int w = 0; w < SCAN_THREADS/32; ++w) tot += s_w[w];
    partial[blockIdx.x] = tot;
  }
}

// one CTA: exclusive scan of the block sums in place, total to partial[n_blocks]
__global__ void __launch_bounds__(1024)
sim_scan_partials_kernel(uint64_t n_blocks, unsigned long long *partial) {
  __shared__ unsigned long long s_w[32];
  __shared__ unsigned long long s_carry;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  if (t == 0) s_carry = 0;
  __syncthreads();
  for (uint64_t b0 = 0; b0 < n_blocks; b0 += 1024) {
    const uint64_t i = b0 + t;
    const unsigned long long v = i < n_blocks ? partial[i] : 0;
    unsigned long long incl = v;
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long up = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += up;
    }
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      unsigned long long w = s_w[lane], wi = w;
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long up = __shfl_up_sync(0xffffffffu, wi, o);
        if (lane >= o) wi += up;
      }
      s_w[lane] = wi - w;                          // exclusive over warps
    }
    __syncthreads();
    const unsigned long long carry = s_carry;
    if (i < n_blocks) partial[i] = carry + s_w[warp] + incl - v;
    __syncthreads();
    if (t == 1023) s_carry = carry + s_w[warp] + incl;
    __syncthreads();
  }
  if (t == 0) partial[n_blocks] = s_carry;
}

__global__ void __launch_bounds__(SCAN_THREADS)
sim_scatter_kernel(uint64_t n, const uint32_t *cnt, const unsigned long long *partial, const double *jbuf, int K,
                   unsigned long long before, unsigned long long *row_off, double *jumps) {
  __shared__ unsigned long long s_w[SCAN_THREADS/32];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const uint64_t first = ((uint64_t)blockIdx.x*SCAN_THREADS + t)*SCAN_ITEMS;
  uint32_t c[SCAN_ITEMS];
  unsigned long long mine = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) { c[k] = first + k < n ? cnt[first + k] : 0; mine += c[k]; }
  unsigned long long incl = mine;
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned long long up = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += up;
  }
  if (lane == 31) s_w[warp] = incl;
  __syncthreads();
  unsigned long long wbase = 0;
  for (int w = 0; w < warp; ++w) wbase += s_w[w];
  unsigned long long at = before + partial[blockIdx.x] + wbase + incl - mine;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    const uint64_t s = first + k;
    if (s < n) {
      row_off[s] = at;
      for (uint32_t j = 0; j < c[k]; ++j) jumps[at + j] = jbuf[s*(uint64_t)K + j];
      at += c[k];
    }
  }
}

struct SimResult {          // kept on the device between epv_simulate and epv_simulate_fetch
  uint64_t n_sites = 0; int n_nodes = 0;
  uint8_t *init = nullptr; unsigned long long *off = nullptr; double *jumps = nullptr;
  uint64_t n_jumps = 0, jumps_cap = 0;
  void release() { cudaFree(init); cudaFree(off); cudaFree(jumps); init = nullptr; off = nullptr; jumps = nullptr; n_sites = 0; }
};

SimResult &result_of(epv_ctx *c) {
  if (!c->sim) c->sim = new SimResult();
  return *static_cast<SimResult *>(c->sim);
}

// sample_state_sequence (EpiEvoModel.cpp:282-298) with a 64-bit mixer as the generator
inline double next_unit(uint64_t &state) {
  uint64_t z = (state += 0x9e3779b97f4a7c15ull);
  z = (z ^ (z >> 30))*0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27))*0x94d049bb133111ebull;
  z ^= z >> 31;
  return ((double)(z >> 11) + 0.5)*(1.0/9007199254740992.0);
}

}  // namespace

void sim_release(epv_ctx *c) {
  if (!c->sim) return;
  SimResult *r = static_cast<SimResult *>(c->sim);
  r->release();
  delete r;
  c->sim = nullptr;
}

}  // namespace epv

using namespace epv;

extern "C" {

int epv_simulate(epv_ctx *c, uint64_t n_sites, const uint8_t *root_states, uint64_t seed,
                 uint64_t *n_jumps_out) {
  if (!c) return 1;
  if (!c->n_nodes) return fail(c, "epv_set_tree must be called before epv_simulate");
  if (!c->have_model) return fail(c, "epv_set_model must be called before epv_simulate");
  if (n_sites < 3) return fail(c, "need at least 3 sites");
  cudaSetDevice(c->device);
  const int n_nodes = c->n_nodes;
  const uint64_t n = n_sites;
  SimRates R;
  R.lambda = 0.0;
  for (int i = 0; i < 8; ++i) { R.rate[i] = c->mp.rate[i]; if (R.rate[i] > R.lambda) R.lambda = R.rate[i]; }
  if (!(R.lambda > 0.0)) return fail(c, "all triplet rates are zero");
  R.inv_lambda = 1.0/R.lambda;
  double t_max = 0.0, t_sum = 0.0;
  for (int v = 1; v < n_nodes; ++v) { t_max = std::max(t_max, c->h_branch[v]); t_sum += c->h_branch[v]; }

  // root sequence
  std::vector<uint8_t> root(n);
  if (root_states) std::memcpy(root.data(), root_states, n);
  else {
    uint64_t st = seed ^ 0x6a09e667f3bcc909ull;
    const double T00 = c->mp.T[0], T11 = c->mp.T[3];
    const double pi1 = (1.0 - T00)/(2.0 - T11 - T00);
    root[0] = next_unit(st) < pi1;
    for (uint64_t i = 1; i < n; ++i) {
      const double r = next_unit(st), p = root[i - 1] ? T11 : T00;
      root[i] = r <= p ? root[i - 1] : (uint8_t)!root[i - 1];
    }
  }

  SimResult &res = result_of(c);
  res.release();
  res.n_sites = n; res.n_nodes = n_nodes;
  const uint64_t n_rows = (uint64_t)n_nodes*n;
  uint8_t *d_end = nullptr;
  unsigned long long *d_pk = nullptr, *d_partial = nullptr;
  uint32_t *d_kc = nullptr, *d_cnt = nullptr;
  double *d_jbuf = nullptr;
  unsigned *d_flags = nullptr;                      // [0] active, [1] error
  auto cleanup = [&]() {
    cudaFree(d_end); cudaFree(d_pk); cudaFree(d_partial); cudaFree(d_kc); cudaFree(d_cnt); cudaFree(d_jbuf);
    cudaFree(d_flags);
  };
  const uint64_t n_scan_blocks = (n + SCAN_THREADS*SCAN_ITEMS - 1)/(SCAN_THREADS*SCAN_ITEMS);
#define SIM_CUDA(call) do { if (int _r = check_cuda(c, (call), #call)) { cleanup(); res.release(); return _r; } } while (0)
  SIM_CUDA(cudaMalloc(&res.init, n_rows));
  SIM_CUDA(cudaMalloc(&res.off, (n_rows + 1)*sizeof(unsigned long long)));
  res.jumps_cap = (uint64_t)(1.5*(double)n*t_sum) + (1u << 20);
  SIM_CUDA(cudaMalloc(&res.jumps, res.jumps_cap*sizeof(double)));
  SIM_CUDA(cudaMalloc(&d_end, n_rows));
  SIM_CUDA(cudaMalloc(&d_pk, n*sizeof(unsigned long long)));
  SIM_CUDA(cudaMalloc(&d_kc, n*sizeof(uint32_t)));
  SIM_CUDA(cudaMalloc(&d_cnt, n*sizeof(uint32_t)));
  SIM_CUDA(cudaMalloc(&d_partial, (n_scan_blocks + 1)*sizeof(unsigned long long)));
  SIM_CUDA(cudaMalloc(&d_flags, 2*sizeof(unsigned)));
  // per-site jump buffer: generous for the expected count, doubled if a row overflows
  int K = (int)std::ceil(8.0 + 3.0*R.lambda*t_max);
  SIM_CUDA(cudaMalloc(&d_jbuf, (size_t)n*K*sizeof(double)));
  cudaStream_t st = c->stream;
  SIM_CUDA(cudaMemcpyAsync(d_end, root.data(), n, cudaMemcpyHostToDevice, st));      // "end state" of the root
  SIM_CUDA(cudaMemsetAsync(res.init, 0, n, st));
  SIM_CUDA(cudaMemsetAsync(res.off, 0, n*sizeof(unsigned long long), st));              // root rows are empty
  const unsigned grid1 = (unsigned)((n + 255)/256), grid_tiles = (unsigned)((n + SIM_TILE - 1)/SIM_TILE);
  uint64_t total = 0;
  for (int v = 1; v < n_nodes; ++v) {
    const int p = c->h_parent[v];
    const double T = c->h_branch[v];
    for (;;) {                                       // repeated only when a jump buffer overflowed
      SIM_CUDA(cudaMemsetAsync(d_flags, 0, 2*sizeof(unsigned), st));
      sim_init_kernel<<<grid1, 256, 0, st>>>(n, 0, v, d_end + (uint64_t)p*n, res.init + (uint64_t)v*n, d_pk, d_kc,
                                             d_cnt, seed, R.inv_lambda, T);
      ++c->launches;
      unsigned flags[2] = {1u, 0u};
      while (flags[0]) {
        SIM_CUDA(cudaMemsetAsync(d_flags, 0, sizeof(unsigned), st));
        sim_round_kernel<<<grid_tiles, SIM_TILE, 0, st>>>(n, 0, v, d_pk, d_kc, d_cnt, d_jbuf, K, R, T, seed,
                                                          d_flags, d_flags + 1);
        ++c->launches;
        SIM_CUDA(cudaMemcpyAsync(flags, d_flags, sizeof(flags), cudaMemcpyDeviceToHost, st));
        SIM_CUDA(cudaStreamSynchronize(st));
      }
      if (!(flags[1] & SIM_ERR_CAPACITY)) break;
      if (K > (1 << 20)) { cleanup(); res.release(); return fail(c, "epv_simulate: a row holds more than 2^20 jumps"); }
      K *= 2;
      cudaFree(d_jbuf); d_jbuf = nullptr;
      SIM_CUDA(cudaMalloc(&d_jbuf, (size_t)n*K*sizeof(double)));
    }
    sim_end_kernel<<<grid1, 256, 0, st>>>(n, d_pk, d_end + (uint64_t)v*n);
    sim_block_sums_kernel<<<(unsigned)n_scan_blocks, SCAN_THREADS, 0, st>>>(n, d_cnt, d_partial);
    sim_scan_partials_kernel<<<1, 1024, 0, st>>>(n_scan_blocks, d_partial);
    c->launches += 3;
    unsigned long long branch_total = 0;
    SIM_CUDA(cudaMemcpyAsync(&branch_total, d_partial + n_scan_blocks, sizeof(branch_total), cudaMemcpyDeviceToHost, st));
    SIM_CUDA(cudaStreamSynchronize(st));
    if (total + branch_total > res.jumps_cap) {      // grow, keeping what is there
      const uint64_t cap = (uint64_t)(1.5*(double)(total + branch_total)*(double)(n_nodes - 1)/(double)v) + (1u << 20);
      double *bigger = nullptr;
      SIM_CUDA(cudaMalloc(&bigger, cap*sizeof(double)));
      SIM_CUDA(cudaMemcpyAsync(bigger, res.jumps, total*sizeof(double), cudaMemcpyDeviceToDevice, st));
      SIM_CUDA(cudaStreamSynchronize(st));
      cudaFree(res.jumps); res.jumps = bigger; res.jumps_cap = cap;
    }
    sim_scatter_kernel<<<(unsigned)n_scan_blocks, SCAN_THREADS, 0, st>>>(n, d_cnt, d_partial, d_jbuf, K, total,
                                                                         res.off + (uint64_t)v*n, res.jumps);
    ++c->launches;
    total += branch_total;
  }
  const unsigned long long grand = total;
  SIM_CUDA(cudaMemcpyAsync(res.off + n_rows, &grand, sizeof(grand), cudaMemcpyHostToDevice, st));
  SIM_CUDA(cudaStreamSynchronize(st));
  SIM_CUDA(cudaGetLastError());
#undef SIM_CUDA
  cleanup();
  res.n_jumps = total;
  if (n_jumps_out) *n_jumps_out = total;
  return 0;
}

int epv_simulate_fetch(epv_ctx *c, uint8_t *init_state, uint64_t *row_off, double *jumps) {
  if (!c) return 1;
  SimResult &res = result_of(c);
  if (!res.n_sites) return fail(c, "epv_simulate_fetch needs a preceding epv_simulate");
  cudaSetDevice(c->device);
  const uint64_t n_rows = (uint64_t)res.n_nodes*res.n_sites;
  EPV_CUDA(c, cudaMemcpyAsync(init_state, res.init, n_rows, cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaMemcpyAsync(row_off, res.off, (n_rows + 1)*sizeof(uint64_t), cudaMemcpyDeviceToHost, c->stream));
  if (res.n_jumps)
    EPV_CUDA(c, cudaMemcpyAsync(jumps, res.jumps, res.n_jumps*sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  res.release();
  return 0;
}

/* The simulated histories become the context's paths WITHOUT leaving the device (for shards too
 * large to stage through host arrays: 10^7 sites x 62 branches is 6 GB of offsets alone).  Same
 * result as epv_simulate_fetch + epv_upload_paths. */
int epv_simulate_adopt(epv_ctx *c) {
  if (!c) return 1;
  SimResult &res = result_of(c);
  if (!res.n_sites) return fail(c, "epv_simulate_adopt needs a preceding epv_simulate");
  const uint64_t n = res.n_sites;
  if (int r = check_upload(c, n, (uint64_t)res.n_nodes, c->h_branch.data())) return r;
  const int n_br = c->n_nodes - 1;
  const uint64_t own_lo = c->halo_lo, own_hi = n - c->halo_hi, n_own = own_hi - own_lo;
  // where every branch's owned rows start and end in the simulator's jump array
  std::vector<unsigned long long> lo(n_br), hi(n_br);
  EPV_CUDA(c, cudaMemcpy2D(lo.data(), 8, res.off + n + own_lo, n*8, 8, (size_t)n_br, cudaMemcpyDeviceToHost));
  EPV_CUDA(c, cudaMemcpy2D(hi.data(), 8, res.off + n + own_hi, n*8, 8, (size_t)n_br, cudaMemcpyDeviceToHost));
  std::vector<uint64_t> count(n_br);
  for (int bi = 0; bi < n_br; ++bi) count[bi] = hi[bi] - lo[bi];
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  if (int r = ensure_shape(c, n)) return r;
  if (int r = ensure_jumps(c, count)) return r;
  c->cur = 0; c->tri_valid = false; c->stats_valid = false;
  launch_import(res.init, res.off, n, c->n_nodes, own_lo, n_own, c->d_off[0], c->off_pitch, c->d_state, c->W, c->stream);
  ++c->launches;
  for (int bi = 0; bi < n_br; ++bi)
    if (count[bi])
      EPV_CUDA(c, cudaMemcpyAsync(c->d_jmp[0] + c->h_jbase[bi], res.jumps + lo[bi], count[bi]*sizeof(double),
                                  cudaMemcpyDeviceToDevice, c->stream));
  // halo rows (<= 6 sites): offsets and jump times through the host into the halo store
  if (c->halo_lo || c->halo_hi) {
    std::vector<unsigned long long> ho((size_t)2*HALO_W*n_br*2, 0);
    uint64_t kmax = 0;
    for (int k = 0; k < 2*HALO_W; ++k) {
      const int64_t s = slot_site(c, k);
      if (s < 0) continue;
      EPV_CUDA(c, cudaMemcpy2D(&ho[(size_t)k*n_br*2], 16, res.off + n + (uint64_t)s, n*8, 16, (size_t)n_br, cudaMemcpyDeviceToHost));
      for (int bi = 0; bi < n_br; ++bi) kmax = std::max<uint64_t>(kmax, ho[((size_t)k*n_br + bi)*2 + 1] - ho[((size_t)k*n_br + bi)*2]);
    }
    // the owned rows next to the ends count too (they are the ones packed later)
    if (kmax > 30000) return fail(c, "a shard-boundary row holds more than 30000 jumps");
    if (int r = set_halo_capacity(c, std::max<int>(HALO_K_MIN, 2*(int)kmax + 16))) return r;
    const size_t HK = c->halo_k;
    std::vector<uint16_t> hc((size_t)2*HALO_W*n_br, 0);
    std::vector<double> hj((size_t)2*HALO_W*n_br*HK, 0.0);
    for (int k = 0; k < 2*HALO_W; ++k) {
      if (slot_site(c, k) < 0) continue;
      for (int bi = 0; bi < n_br; ++bi) {
        const unsigned long long a = ho[((size_t)k*n_br + bi)*2], b = ho[((size_t)k*n_br + bi)*2 + 1];
        hc[(size_t)k*n_br + bi] = (uint16_t)(b - a);
        if (b > a) EPV_CUDA(c, cudaMemcpy(&hj[((size_t)k*n_br + bi)*HK], res.jumps + a, (b - a)*sizeof(double), cudaMemcpyDeviceToHost));
      }
    }
    EPV_CUDA(c, cudaMemcpyAsync(c->d_hcnt, hc.data(), hc.size()*sizeof(uint16_t), cudaMemcpyHostToDevice, c->stream));
    EPV_CUDA(c, cudaMemcpyAsync(c->d_hjmp, hj.data(), hj.size()*sizeof(double), cudaMemcpyHostToDevice, c->stream));
  }
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  EPV_CUDA(c, cudaGetLastError());
  res.release();
  return 0;
}

}  // extern "C"
