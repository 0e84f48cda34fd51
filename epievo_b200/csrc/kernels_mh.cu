// kernels_mh.cu -- one stride-3 colour pass of the Metropolis-Hastings sweep
// (SingleSiteSampler::single_iteration, SingleSiteSampler.cpp:538-548, re-ordered into
// stride-3 colour classes: updates of s and s+3 touch disjoint data), as TWO kernels:
//
//   mh_update_kernel    one thread per group of 3 sites runs the update of the group's
//                       site of the pass's colour (mh_site_flat).  State bits and tri_llh
//                       are updated in place; an accepted proposal's rows go to a
//                       staging blob (u16 jump counts per branch + jump times), found
//                       through stage_off[group].
//   csr_rebuild_kernel  re-compacts the per-branch CSR into the ping-pong partner:
//                       per branch, new row lengths -> tile scan -> single-pass chained
//                       scan (decoupled look-back) across tiles -> offsets and jump
//                       times written in site order.  Pure streaming, every tile costs
//                       the same, so the look-back never waits for long.
//
// The two were one kernel at first.  ncu showed a third of its instructions in the
// look-back spin: tiles whose updates finished early waited, holding their SM slots, for
// the slowest earlier tile (update time varies with the rows' jump counts).  Splitting
// costs one staging round trip of the accepted proposals (a few % of the row bytes).
#include <cstdlib>
#include <type_traits>

#include "kernels.h"
#include "mh_flat.cuh"
#include "mh_lock.cuh"
#include "mh_mix.cuh"
#include "scan.cuh"

namespace epv {

namespace {

constexpr unsigned MH_OVF_BLOCKS = 296;     // grid of the overflow launch (grid-stride over the list)

// The main update launch defers the sites its flat routine cannot finish to a second launch
// (see mh_update_group).  Measured ms per pass, deferred vs inline fallback: 8-leaf tree 5.18 vs
// 5.48, 4-branch test tree 5.64 vs 6.45, 32-leaf tree 8.73 vs 7.54 -- so only up to 16 nodes.
__host__ __device__ constexpr bool mh_defer(int maxn) { return maxn <= 16; }

__device__ __forceinline__ int stage_header_words(int n_br) { return (n_br + 3) >> 2; }

// CTAs per SM the update kernel is compiled for (register budget 65536 / (threads x this)).
// Measured ms per pass at 3 / 4 / 5 CTAs of 128 threads: 8-leaf tree (MAXN 16) 6.1 / 5.5 / 6.5; 4-branch
// test tree (MAXN 8) 6.5 / 6.5 / 7.4; 32-leaf tree (MAXN 64) 7.7 / 8.9 / 10.6.
//
// Threads per CTA.  The halo-free main launch runs ONE large CTA per SM instead of 4 (3) CTAs of 128
// threads: the 16 (12) warps of an SM then start their site updates together and walk the ~190 KB routine
// roughly in phase, so they share the instruction cache -- with four CTAs at four different places of the
// routine, instruction fetch was the largest stall of the test tree's pass (no_instruction 3.1 cycles per
// issue against 0.3 with one CTA of 512, ncu profiles/r02_update_cta_size_ab.txt) and the pass took 5.60
// instead of 4.48 ms (8-leaf tree 5.12 -> 4.91, 32-leaf tree 7.49 -> 7.29 at 512; CTAs of 64 / 32 threads:
// 5.24 / 5.33 on the 8-leaf tree).  The small launches (halo-aware edge groups, overflow list) and the
// lockstep / mix build variants (shared arrays with a 128-column stride) keep 128.
#ifndef EPV_MH_T16
#define EPV_MH_T16 512        // main launch, trees up to 16 nodes
#endif
#ifndef EPV_MH_T64
#define EPV_MH_T64 512        // main launch, trees up to 64 nodes (384 x 1 at 156 registers measured slower: 7.64 vs 7.28 ms)
#endif
__host__ __device__ constexpr int mh_impl(int maxn);
__host__ __device__ constexpr int mh_threads(int maxn, bool halo) {
  return (halo || maxn > 64 || mh_impl(maxn) != 0) ? MH_THREADS : (maxn <= 16 ? EPV_MH_T16 : EPV_MH_T64);
}
#ifdef EPV_MH_MINBLOCKS
__host__ __device__ constexpr int mh_minblocks(int, bool) { return EPV_MH_MINBLOCKS; }
#else
__host__ __device__ constexpr int mh_minblocks(int maxn, bool halo) {
  return (maxn <= 16 ? 512 : 384)/mh_threads(maxn, halo) > 0 ? (maxn <= 16 ? 512 : 384)/mh_threads(maxn, halo) : 1;
}
#endif

// The lockstep routine (mh_lock.cuh: branch loop warp-uniform, table-driven single-interval branches)
// reproduces the flat routine's chains bit for bit (tests/test_gpu_chain_golden.py) but measured SLOWER
// on the B200: 6.56 vs 5.12 ms per pass (8-leaf tree), 7.30 vs 5.64 (test tree), 8.48 vs 7.52 (32-leaf
// tree), 0.44 vs 0.41 (1e6-site single branch) -- the interval loop of the lanes that do have neighbour
// jumps runs per BRANCH (max over lanes of the branch's intervals, ~2.2 heavy iterations per branch at 2-4
// active lanes) instead of once per site.  Kept behind EPV_MH_LOCK as the base of the next step.
// EPV_MH_IMPL: 0 = mh_site_flat, 1 = mh_site_lock, 2 = mh_site_mix (mh_mix.cuh: lockstep for the
// single-interval branches, per-lane flat loops for the others); trees wider than 64 nodes always 0.
// Measured on the B200 (ms per colour pass, flat / lock / mix): 8-leaf tree 5.12 / 6.56 / 8.27, test tree
// 5.64 / 7.30 / 8.33, 32-leaf tree 7.52 / 8.48 / 19.2 (its shared arrays leave 2 CTAs per SM), 1e6-site
// single branch 0.41 / 0.44 / 0.53: the flat routine stays the default; all three pass the exact-posterior
// tests (tests/test_gpu_exact_posterior.py).
#ifndef EPV_MH_IMPL
#define EPV_MH_IMPL 0
#endif
#ifndef EPV_MH_QSMEM            // lock / mix: Felsenstein q of the internal nodes in shared memory (else local)
#define EPV_MH_QSMEM 1
#endif
__host__ __device__ constexpr int mh_impl(int maxn) { return maxn <= 64 ? EPV_MH_IMPL : 0; }
__host__ __device__ constexpr bool mh_lock(int maxn) { return mh_impl(maxn) != 0; }

// jump-queue slots per thread (shared memory: 24 B each)
__host__ __device__ constexpr int mh_qcap(int maxn) { return maxn <= 16 ? 8 : maxn <= 64 ? 16 : 24; }

// The update of group g's site of the pass's colour, and the staging of its rows.
// DEFER (the halo-free main launch of trees up to 16 nodes): the flat routine is tried once; a site whose proposal does
// not fit the per-thread buffer (-1) or whose neighbours' intervals do not fit the per-thread
// stack (-3) is appended to the overflow list and left to a second, small launch of the full
// routine below.  The main kernel then carries one copy of the flat routine and nothing else
// (it was 190 KB of SASS with the retry and the generic fallback inlined; instruction fetch was
// 14 % of its stall cycles).  !DEFER: the full routine -- flat, then the generic per-branch code,
// then a rerun with a global scratch slot -- all decisions depend on the site's data only, so a
// site gets the same result whichever launch takes it.
// Shared memory of the update kernel, in this order: jump queue [THREADS*QCAP] QSlot, prefix counts
// [THREADS] int, then (lockstep routine, trees up to 64 nodes) TreeTab [n_nodes], BranchTab [n_nodes],
// jump counts per node u16 [n_nodes][THREADS], q of the internal nodes f64 [2*slots][THREADS].
#ifndef EPV_JBASE_SMEM
#define EPV_JBASE_SMEM 0
#endif
struct MhShared {
  QSlot *queue; int *excl; TreeTab *tt; BranchTab *bt; uint16_t *pn; double *q; uint16_t *poff;
  const uint64_t *jbase;
};
template <int MAXN, int THREADS>
__host__ __device__ inline size_t mh_smem_layout(int n_nodes, int q_slots, size_t *off /*[7]*/) {
  size_t at = 0;
  off[0] = at; at += (size_t)THREADS*mh_qcap(MAXN)*sizeof(QSlot);
  off[1] = at; at += (size_t)THREADS*sizeof(int);
  if (mh_lock(MAXN)) {
    at = (at + 15) & ~(size_t)15;
    off[2] = at; at += (size_t)n_nodes*sizeof(TreeTab);
    off[3] = at; at += (size_t)n_nodes*sizeof(BranchTab);
    off[4] = at; at += (size_t)n_nodes*THREADS*sizeof(uint16_t);
    at = (at + 15) & ~(size_t)15;
    off[5] = at; at += (size_t)2*q_slots*THREADS*sizeof(double);
    off[6] = at; if (mh_impl(MAXN) == 2) at += (size_t)n_nodes*THREADS*sizeof(uint16_t);
  }
  else off[2] = off[3] = off[4] = off[5] = off[6] = at;
  return at;
}

template <int MAXN, int MAXJ, bool HALO, bool DEFER, bool QSMEM>
__device__ __forceinline__ void
mh_update_group(const PassArgs &a, const ModelParams &s_mp, const CtxTables &s_tb, const MhShared &sh, uint64_t g) {
  const int t = threadIdx.x, lane = t & 31;
  const int n_br = a.tree.n_nodes - 1;
  const uint64_t st = 3*g + a.colour;                          // the group's target site
  const int64_t gsite = a.global0 + (int64_t)st;
  constexpr bool LOCK = mh_lock(MAXN);

  // interval stack / jump queue capacities grow with the tree (expected use ~ 0.1 / 0.07 per branch)
  constexpr int PB = MAXN <= 16 ? 16 : MAXN/2, QCAP = mh_qcap(MAXN);
  double pj_local[MAXJ];
  uint16_t pn[MAXN];                   // generic routine / wide trees only
  uint8_t pend[MAXN];
  Proposal pr; pr.pj = pj_local; pr.cap = MAXJ; pr.pn = pn; pr.pend = pend;
  int accepted = 0, total = 0;
  bool deferred = false, local_pn = !LOCK;
  typename LockMask<MAXN <= 64 ? MAXN : 64>::type propmask = 0, pend_bits = 0;
  const bool active = g < a.n_groups && st >= a.upd_lo && st < a.upd_hi && gsite >= 1 &&
                      gsite + 1 < (int64_t)a.n_global;
  SiteCtx c; c.pv = a.in; c.state = a.state; c.W = a.W; c.tree = a.tree; c.site = st;
#if EPV_JBASE_SMEM
  c.pv.jbase = sh.jbase;                 // the branches' slice starts from shared memory (a dependent load per row access)
#endif
  constexpr bool MIX = mh_impl(MAXN) == 2;
  if constexpr (LOCK) {
    typename std::conditional<MIX, MixScratch<MAXN, PB, QCAP, QSMEM>, LockScratch<MAXN, PB, QCAP, QSMEM> >::type S;
    S.queue = sh.queue + t*QCAP; S.excl = sh.excl + (t & ~31);
    S.pn_s = sh.pn + t; S.q_s = sh.q + t;
    if constexpr (MIX) S.poff_s = sh.poff + t;
    auto site_update = [&](bool act, double *pj, int cap, auto &pe, auto &pm, int &tot) {
      if constexpr (MIX)
        return mh_site_mix<MAXN, PB, QCAP, HALO, QSMEM>(act, a.coop != 0, s_mp, s_tb, sh.tt, sh.bt, c, a.state, a.tri_llh,
                                                        a.seed, a.sweep, (uint64_t)gsite, a.n_global, S, pj, cap, pe, pm, tot);
      else
        return mh_site_lock<MAXN, PB, QCAP, HALO, QSMEM>(act, a.coop != 0, s_mp, s_tb, sh.tt, sh.bt, c, a.state, a.tri_llh,
                                                         a.seed, a.sweep, (uint64_t)gsite, a.n_global, S, pj, cap, pe, pm, tot);
    };
    int r = site_update(active, pj_local, MAXJ, pend_bits, propmask, total);
    if (DEFER) {
      if (r == -1 || r == -3) {
        a.ovf_list[1 + atomicAdd(a.ovf_list, 1u)] = (uint32_t)g;
        deferred = true; r = 0;
      }
    }
    else {
      // -3: interval stack too small -- a property of the neighbours' rows only, so switching
      // to the generic per-branch routine here is unbiased (rare; runs divergently)
      const bool generic = r == -3;
      if (generic) {
        r = mh_site(s_mp, c, a.state, a.tri_llh, a.seed, a.sweep, (uint64_t)gsite, a.n_global,
                    S.q_local, pr, total);
        local_pn = true;
      }
      // -1: proposal longer than the per-thread buffer: rerun with a global scratch slot and
      // the same streams (=> the same proposal).  The lockstep routine is warp-synchronous, so the
      // whole warp re-enters it with only the overflowing lanes active.
      if (__any_sync(0xffffffffu, r == -1)) {
        const bool again = r == -1;
        if (again) {
          const uint32_t slot = atomicAdd(a.scratch_ctr, 1u);
          if (slot >= a.scratch_slots) { atomicOr(a.err, ERR_SCRATCH_SLOTS); r = 0; }
          else { pr.pj = a.scratch + (uint64_t)slot*a.scratch_cap; pr.cap = (int)a.scratch_cap; }
        }
        const bool redo = again && r == -1;
        typename LockMask<MAXN>::type pm2 = 0, pe2 = 0; int tot2 = 0;
        const int r2 = site_update(redo && !generic, pr.pj, pr.cap, pe2, pm2, tot2);
        if (redo && !generic) { r = r2; propmask = pm2; pend_bits = pe2; total = tot2; }
        if (redo && generic)
          r = mh_site(s_mp, c, a.state, a.tri_llh, a.seed, a.sweep, (uint64_t)gsite, a.n_global,
                      S.q_local, pr, total);
        if (r == -1) { atomicOr(a.err, ERR_SCRATCH_SIZE); r = 0; }
      }
    }
    if (r == -2) { atomicOr(a.err, ERR_SAMPLER_STUCK); r = 0; }
    if (r < 0) r = 0;
    accepted = r;
  }
  else {
    FlatScratch<MAXN, PB, QCAP> S;
    S.queue = sh.queue + t*QCAP; S.excl = sh.excl + (t & ~31);
    int r = mh_site_flat<MAXN, PB, QCAP, HALO>(active, a.coop != 0, s_mp, s_tb, c, a.state, a.tri_llh, a.seed, a.sweep,
                                               (uint64_t)gsite, a.n_global, S, pr, total);
    if (DEFER) {
      // the main launch of a long pass: a site the flat routine cannot finish goes to the overflow list
      // (second, small launch of the full routine below); the main kernel carries no fallback code
      if (r == -1 || r == -3) {
        a.ovf_list[1 + atomicAdd(a.ovf_list, 1u)] = (uint32_t)g;
        deferred = true; r = 0;
      }
    }
    const bool generic = !DEFER && r == -3;
    if (generic)
      r = mh_site(s_mp, c, a.state, a.tri_llh, a.seed, a.sweep, (uint64_t)gsite, a.n_global, S.q, pr, total);
    if (!DEFER && __any_sync(0xffffffffu, r == -1)) {
      const bool again = r == -1;
      if (again) {
        const uint32_t slot = atomicAdd(a.scratch_ctr, 1u);
        if (slot >= a.scratch_slots) { atomicOr(a.err, ERR_SCRATCH_SLOTS); r = 0; }
        else { pr.pj = a.scratch + (uint64_t)slot*a.scratch_cap; pr.cap = (int)a.scratch_cap; }
      }
      const bool redo = again && r == -1;
      const int r2 = mh_site_flat<MAXN, PB, QCAP, HALO>(redo && !generic, a.coop != 0, s_mp, s_tb, c, a.state, a.tri_llh,
                                                        a.seed, a.sweep, (uint64_t)gsite, a.n_global, S, pr, total);
      if (redo && !generic) r = r2;
      if (redo && generic)
        r = mh_site(s_mp, c, a.state, a.tri_llh, a.seed, a.sweep, (uint64_t)gsite, a.n_global, S.q, pr, total);
      if (r == -1) { atomicOr(a.err, ERR_SCRATCH_SIZE); r = 0; }
    }
    if (r == -2) { atomicOr(a.err, ERR_SAMPLER_STUCK); r = 0; }
    if (r < 0) r = 0;
    accepted = r;
  }
  const unsigned acc_mask = __ballot_sync(0xffffffffu, accepted);
  if (lane == 0 && acc_mask) atomicAdd(a.n_accept, (unsigned long long)__popc(acc_mask));

  // ---- stage the accepted proposal's rows -----------------------------------------
  if (g >= a.n_groups || deferred) return;
  uint32_t so = STG_REJECTED;
  if (accepted) {
    if (total == 0) so = STG_EMPTY;
    else {
      const int hw = stage_header_words(n_br);
      const unsigned long long need = (unsigned long long)hw + total;
      const unsigned long long at = atomicAdd(a.stage_ctr, need);
      if (at + need > a.stage_cap) atomicOr(a.err, ERR_STAGE_CAPACITY);
      else {
        so = (uint32_t)at;
        double *blob = a.stage + at;
        uint16_t *cnt = reinterpret_cast<uint16_t *>(blob);
        int k = 0;
        for (int bi = 0; bi < 4*hw; ++bi) {
          uint16_t n = 0;
          if (bi < n_br) n = local_pn ? pn[bi + 1] : (((propmask >> (bi + 1)) & 1) ? sh.pn[(bi + 1)*mh_threads(MAXN, HALO) + t] : 0);
          cnt[bi] = n;
          if (MIX && !local_pn && n) {                 // rows are not in branch order in pj: gather
            const double *src = pr.pj + sh.poff[(bi + 1)*mh_threads(MAXN, HALO) + t];
            for (int x = 0; x < n; ++x) blob[hw + k++] = src[x];
          }
        }
        if (!MIX || local_pn) for (k = 0; k < total; ++k) blob[hw + k] = pr.pj[k];
      }
    }
  }
  a.stage_off[g] = so;
}

// HALO = false: the main launch (one site per thread, failures deferred).  HALO = true: the full
// routine on (a) the few groups next to a shard's ends, whose neighbours live in the halo store,
// and (b) the overflow list of the main launch (a.list_mode; grid-stride over the list, whose
// length is only known on the device).
template <int MAXN, int MAXJ, bool HALO, bool DEFER, bool QSMEM>
__global__ void __launch_bounds__(mh_threads(MAXN, HALO), mh_minblocks(MAXN, HALO))
mh_update_kernel(const PassArgs a, int q_slots) {
  constexpr int MT = mh_threads(MAXN, HALO);
  __shared__ ModelParams s_mp;
  __shared__ CtxTables s_tb;
  extern __shared__ __align__(16) unsigned char s_dyn[];
  const int t = threadIdx.x;
#if EPV_JBASE_SMEM
  __shared__ uint64_t s_jbase[MAXN];
  for (int i = t; i < a.tree.n_nodes - 1; i += MT) s_jbase[i] = a.in.jbase[i];
#endif
  if (t == 0) s_mp = a.mp;
  fill_tables(a.mp, s_tb, t);
  MhShared sh;
  {
    size_t off[7];
    mh_smem_layout<MAXN, MT>(a.tree.n_nodes, q_slots, off);
    sh.queue = reinterpret_cast<QSlot *>(s_dyn + off[0]); sh.excl = reinterpret_cast<int *>(s_dyn + off[1]);
    sh.tt = reinterpret_cast<TreeTab *>(s_dyn + off[2]); sh.bt = reinterpret_cast<BranchTab *>(s_dyn + off[3]);
    sh.pn = reinterpret_cast<uint16_t *>(s_dyn + off[4]); sh.q = reinterpret_cast<double *>(s_dyn + off[5]);
    sh.poff = reinterpret_cast<uint16_t *>(s_dyn + off[6]);
#if EPV_JBASE_SMEM
    sh.jbase = s_jbase;
#else
    sh.jbase = a.in.jbase;
#endif
  }
  __syncthreads();
  if (mh_lock(MAXN)) {
    fill_lock_tables(s_mp, s_tb, a.tree, sh.tt, sh.bt, t, MT);
    __syncthreads();
  }
  if (!HALO) {
    uint64_t gi = (uint64_t)blockIdx.x*MT + t;
    if (a.reverse && gi < a.g_count) gi = a.g_count - 1 - gi;
    const uint64_t g = gi >= a.g_count ? a.n_groups                // group of sites 3g..3g+2
                                       : a.g_first + gi + (gi >= a.g_split ? a.g_skip : 0);
    mh_update_group<MAXN, MAXJ, false, DEFER, QSMEM>(a, s_mp, s_tb, sh, g);
  }
  else {
    const uint64_t n_items = a.list_mode ? (uint64_t)a.ovf_list[0] : a.g_count;
    const uint64_t stride = (uint64_t)gridDim.x*MT;
    // (trip count uniform within a warp: the routine is warp-synchronous)
    for (uint64_t base = (uint64_t)blockIdx.x*MT; base + (uint64_t)(t & ~31) < n_items; base += stride) {
      const uint64_t gi = base + t;
      uint64_t g = a.n_groups;
      if (gi < n_items) g = a.list_mode ? (uint64_t)a.ovf_list[1 + gi] : a.g_first + gi + (gi >= a.g_split ? a.g_skip : 0);
      mh_update_group<MAXN, MAXJ, true, false, QSMEM>(a, s_mp, s_tb, sh, g);
      __syncwarp();
    }
  }
}

// ---- CSR re-compaction ------------------------------------------------------------------
// Thread t of tile k owns the 4 consecutive CSR rows R = 4*(k*RB_THREADS + t) .. R+3 of
// every branch (offsets move as one 16-byte vector).  Rows whose site has the pass's colour
// take their new length / jump times from the staging blob when the update was accepted.
#ifdef EPV_RB_MINBLOCKS         // CTAs per SM the rebuild kernel is compiled for (default: the compiler's choice)
#define EPV_RB_BOUNDS __launch_bounds__(RB_THREADS, EPV_RB_MINBLOCKS)
#else
#define EPV_RB_BOUNDS __launch_bounds__(RB_THREADS)
#endif
template <int UNROLL>
__global__ void EPV_RB_BOUNDS
csr_rebuild_kernel(const PassArgs a) {
  extern __shared__ uint32_t s_len[];            // [n_branches][RB_THREADS]
  __shared__ uint32_t s_tile;
  __shared__ int s_fail;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int n_br = a.tree.n_nodes - 1;
  if (t == 0) {
    s_tile = atomicAdd(a.ticket, 1u) - a.ticket_base;   // tile order = scheduling order
    s_fail = 0;
  }
  __syncthreads();
  const uint32_t tile = s_tile;
  const uint64_t own_lo = a.in.own_lo, n_own = a.in.own_hi - own_lo;
  const uint64_t R = 4*((uint64_t)tile*RB_THREADS + t);
  const int hw = stage_header_words(n_br);

  // the (at most two) rows of the quad whose site has this pass's colour: row k0 and, when
  // k0 == 0, row 3.  Kept in scalars (no dynamically indexed local arrays).
  const int k0 = (int)((a.colour + 3 - (int)((R + own_lo) % 3)) % 3);
  bool acc0 = false, acc1 = false;
  const uint16_t *cnt0 = nullptr, *cnt1 = nullptr;
  const double *pj0 = nullptr, *pj1 = nullptr;
  if (R + k0 < n_own) {
    const uint32_t so = a.stage_off[(R + k0 + own_lo)/3];
    acc0 = so != STG_REJECTED;
    if (so < STG_EMPTY) { cnt0 = reinterpret_cast<const uint16_t *>(a.stage + so); pj0 = a.stage + so + hw; }
  }
  if (k0 == 0 && R + 3 < n_own) {
    const uint32_t so = a.stage_off[(R + 3 + own_lo)/3];
    acc1 = so != STG_REJECTED;
    if (so < STG_EMPTY) { cnt1 = reinterpret_cast<const uint16_t *>(a.stage + so); pj1 = a.stage + so + hw; }
  }

  // ---- new row lengths of the quad, summed per thread -------------------------------
  // UNROLL branches at a time with all their loads issued first (the kernel is bound by
  // load latency, not by issue slots)
  for (int b0 = 0; b0 < n_br; b0 += UNROLL) {
    uint4 q[UNROLL];
    uint32_t e[UNROLL], c0[UNROLL], c1[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int bi = min(b0 + u, n_br - 1);
      q[u] = make_uint4(0, 0, 0, 0); e[u] = 0; c0[u] = 0; c1[u] = 0;
      if (R < n_own) {
        const uint32_t *o = a.in.off + (uint64_t)bi*a.in.off_pitch + R;
        q[u] = *reinterpret_cast<const uint4 *>(o);
        e[u] = o[4];
        if (cnt0) c0[u] = cnt0[bi];
        if (cnt1) c1[u] = cnt1[bi];
      }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int bi = b0 + u;
      if (bi >= n_br) break;
      uint32_t sum = 0;
      if (R < n_own) {
        uint32_t len[4] = {q[u].y - q[u].x, q[u].z - q[u].y, q[u].w - q[u].z, e[u] - q[u].w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (R + k >= n_own) len[k] = 0;
          if (k < 3 && k == k0 && acc0) len[k] = c0[u];
          if (k == 3 && acc1) len[k] = c1[u];
          sum += len[k];
        }
      }
      s_len[bi*RB_THREADS + t] = sum;
    }
  }
  __syncthreads();

  // ---- per-branch tile scan + chained look-back -------------------------------------
  constexpr int PER = RB_THREADS/32;
  for (int bi = warp; bi < n_br; bi += PER) {
    uint32_t *row = s_len + bi*RB_THREADS;
    uint32_t v[PER], mine = 0;
#pragma unroll
    for (int k = 0; k < PER; ++k) { v[k] = row[lane*PER + k]; mine += v[k]; }
    uint32_t incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += up;
    }
    const uint32_t tot = __shfl_sync(0xffffffffu, incl, 31);
    const uint32_t base = tile_base(a.desc + (uint64_t)bi*a.n_tiles, tile, a.epoch, tot, lane);
    if (lane == 0 && (uint64_t)base + tot > a.jcap[bi]) {
      // nothing of this tile is written; the scan goes on, so the last tile records the branch's new total
      // and the host can regrow and rerun the rebuild (its inputs are intact)
      atomicOr(a.err, ERR_JUMP_CAPACITY);
      atomicMax(a.jneed + bi, (unsigned long long)base + tot);
      s_fail = 1;
    }
    uint32_t run = base + incl - mine;
#pragma unroll
    for (int k = 0; k < PER; ++k) { row[lane*PER + k] = run; run += v[k]; }
  }
  __syncthreads();
  if (s_fail || R > n_own) return;

  // ---- write the rebuilt rows ------------------------------------------------------------
  // (same grouping: the loads of UNROLL branches are in flight together)
  int poff0 = 0, poff1 = 0;
  for (int b0 = 0; b0 < n_br; b0 += UNROLL) {
    uint4 q[UNROLL];
    uint32_t e[UNROLL], c0[UNROLL], c1[UNROLL];
    uint64_t jb[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int bi = min(b0 + u, n_br - 1);
      q[u] = make_uint4(0, 0, 0, 0); e[u] = 0; c0[u] = 0; c1[u] = 0;
      jb[u] = a.in.jbase[bi];
      if (R < n_own) {
        const uint32_t *o = a.in.off + (uint64_t)bi*a.in.off_pitch + R;
        q[u] = *reinterpret_cast<const uint4 *>(o);
        e[u] = o[4];
        if (cnt0) c0[u] = cnt0[bi];
        if (cnt1) c1[u] = cnt1[bi];
      }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int bi = b0 + u;
      if (bi >= n_br) break;
      const double *src_base = a.in.jmp + jb[u];
      double *dst_base = a.jmp_out + jb[u];
      uint32_t w = s_len[bi*RB_THREADS + t];
      uint32_t out[4] = {w, w, w, w};                  // R == n_own: the closing offset
      if (R < n_own) {
        const uint32_t in[5] = {q[u].x, q[u].y, q[u].z, q[u].w, e[u]};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          out[k] = w;
          if (R + k >= n_own) continue;
          if (k < 3 && k == k0 && acc0) {
            const int n_new = (int)c0[u];
            for (int x = 0; x < n_new; ++x) dst_base[w + x] = pj0[poff0 + x];
            poff0 += n_new;
            w += n_new;
          }
          else if (k == 3 && acc1) {
            const int n_new = (int)c1[u];
            for (int x = 0; x < n_new; ++x) dst_base[w + x] = pj1[poff1 + x];
            poff1 += n_new;
            w += n_new;
          }
          else {
            const uint32_t n = in[k + 1] - in[k];
            for (uint32_t x = 0; x < n; ++x) dst_base[w + x] = src_base[in[k] + x];
            w += n;
          }
        }
      }
      *reinterpret_cast<uint4 *>(a.off_out + (uint64_t)bi*a.in.off_pitch + R) = make_uint4(out[0], out[1], out[2], out[3]);
    }
  }
}

// ---- CSR re-compaction, offsets staged with the bulk-async copy engine ------------------------------------
// Same tiles, same thread -> rows mapping, same chained scan and the same result as csr_rebuild_kernel; what
// changes is how the offsets move.  They are 1.4 GB of the kernel's 1.8 GB of traffic and, per (tile, branch),
// one contiguous 4 KB run: instead of two passes of dependent 16-byte loads per thread (the old kernel waits
// 9.8 cycles per issue on the long scoreboard), one elected thread has the copy engine (cp.async.bulk +
// mbarrier, the TMA path for linear data) bring the runs of a CHUNK of branches into shared memory while the
// CTA works on the previous chunk; the new offsets are written in place in shared memory and go back to HBM
// as bulk stores.  Jump times (4 % of the bytes) are copied as before.
namespace bulk {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned parity) {
  asm volatile(
    "{\n\t.reg .pred p;\n\t"
    "WAIT_%=:\n\t"
    "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
    "@!p bra WAIT_%=;\n\t}"
    :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared, completion counted in bytes on the mbarrier
__device__ __forceinline__ void load_async(void *dst_smem, const void *src_gmem, unsigned bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// shared -> global, tracked by the issuing thread's bulk group
__device__ __forceinline__ void store_async(void *dst_gmem, const void *src_smem, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               :: "l"(dst_gmem), "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void store_wait_read_all() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

constexpr int ROW_ELEMS = RB_TILE_ROWS + 4;       // a tile's offsets of one branch + the closing one, 16-byte multiple

}  // namespace bulk

// Dynamic shared memory: [2][BC][ROW_ELEMS] u32 staged offsets, [BC][RB_THREADS] u32 quad sums, [BC][8] warp-edge
// offsets, 2 mbarriers.
__global__ void __launch_bounds__(RB_THREADS)
csr_rebuild_bulk_kernel(const PassArgs a, int BC) {
  using namespace bulk;
  extern __shared__ __align__(128) unsigned char s_raw[];
  uint32_t *s_off = reinterpret_cast<uint32_t *>(s_raw);                        // [2][BC][ROW_ELEMS]
  uint32_t *s_len = s_off + (size_t)2*BC*ROW_ELEMS;                             // [BC][RB_THREADS]
  uint32_t *s_edge = s_len + (size_t)BC*RB_THREADS;                             // [BC][RB_THREADS/32]
  uint64_t *s_bar = reinterpret_cast<uint64_t *>(s_edge + (size_t)BC*(RB_THREADS/32));   // [2]
  __shared__ uint32_t s_tile;
  __shared__ int s_fail;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int n_br = a.tree.n_nodes - 1;
  if (t == 0) {
    s_tile = atomicAdd(a.ticket, 1u) - a.ticket_base;   // tile order = scheduling order
    s_fail = 0;
    mbar_init(&s_bar[0], 1); mbar_init(&s_bar[1], 1);
    fence_mbar_init();
  }
  __syncthreads();
  const uint32_t tile = s_tile;
  const uint64_t own_lo = a.in.own_lo, n_own = a.in.own_hi - own_lo;
  const uint64_t R0 = (uint64_t)tile*RB_TILE_ROWS, R = R0 + 4*(uint64_t)t;
  const int hw = stage_header_words(n_br);
  const int n_chunks = (n_br + BC - 1)/BC;
  // elements of a branch's offsets to stage (never past the branch's pitch) and to store back
  // (rows < n_own, the closing offset, rounded up to the quad: what the old kernel's threads with R <= n_own write)
  const unsigned load_elems = (unsigned)((a.in.off_pitch - R0) < (uint64_t)ROW_ELEMS ? (a.in.off_pitch - R0) : (uint64_t)ROW_ELEMS);
  const unsigned store_elems = R0 + RB_TILE_ROWS <= n_own ? (unsigned)RB_TILE_ROWS : (unsigned)(((n_own - R0) & ~3ull) + 4);

  auto issue_loads = [&](int chunk) {        // thread 0 only
    const int s = chunk & 1, b0 = chunk*BC, nb = min(BC, n_br - b0);
    mbar_expect_tx(&s_bar[s], (unsigned)nb*load_elems*4u);
    for (int u = 0; u < nb; ++u)
      load_async(s_off + ((size_t)s*BC + u)*ROW_ELEMS, a.in.off + (uint64_t)(b0 + u)*a.in.off_pitch + R0, load_elems*4u, &s_bar[s]);
  };
  if (t == 0) issue_loads(0);


  // the (at most two) rows of the quad whose site has this pass's colour: row k0 and, when k0 == 0, row 3
  const int k0 = (int)((a.colour + 3 - (int)((R + own_lo) % 3)) % 3);
  bool acc0 = false, acc1 = false;
  const uint16_t *cnt0 = nullptr, *cnt1 = nullptr;
  const double *pj0 = nullptr, *pj1 = nullptr;
  if (R + k0 < n_own) {
    const uint32_t so = a.stage_off[(R + k0 + own_lo)/3];
    acc0 = so != STG_REJECTED;
    if (so < STG_EMPTY) { cnt0 = reinterpret_cast<const uint16_t *>(a.stage + so); pj0 = a.stage + so + hw; }
  }
  if (k0 == 0 && R + 3 < n_own) {
    const uint32_t so = a.stage_off[(R + 3 + own_lo)/3];
    acc1 = so != STG_REJECTED;
    if (so < STG_EMPTY) { cnt1 = reinterpret_cast<const uint16_t *>(a.stage + so); pj1 = a.stage + so + hw; }
  }
  int poff0 = 0, poff1 = 0;

  for (int chunk = 0; chunk < n_chunks; ++chunk) {
    const int s = chunk & 1, b0 = chunk*BC, nb = min(BC, n_br - b0);
    // the other stage was stored from by chunk-1: its bulk stores must have READ shared memory before the copy
    // engine overwrites it with chunk+1 (all threads left chunk-1 at the barrier that closed it)
    if (t == 0 && chunk + 1 < n_chunks) { store_wait_read_all(); issue_loads(chunk + 1); }
    // staged jump counts of this chunk's branches (global loads, issued before the wait)
    uint32_t c0[8], c1[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      c0[u] = 0; c1[u] = 0;
      if (u < nb) { if (cnt0) c0[u] = cnt0[b0 + u]; if (cnt1) c1[u] = cnt1[b0 + u]; }
    }
    mbar_wait(&s_bar[s], (unsigned)((chunk >> 1) & 1));
    uint32_t *so_base = s_off + (size_t)s*BC*ROW_ELEMS;

    // ---- new row lengths of the quad, summed per thread ------------------------------------------------
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      if (u >= nb) break;
      uint32_t sum = 0;
      const uint32_t *o = so_base + (size_t)u*ROW_ELEMS + 4*t;
      const uint32_t e = o[4];
      if (lane == 31) s_edge[u*(RB_THREADS/32) + warp] = e;     // the next warp's first offset, kept for the write phase
      if (R < n_own) {
        const uint4 q = *reinterpret_cast<const uint4 *>(o);
        uint32_t len[4] = {q.y - q.x, q.z - q.y, q.w - q.z, e - q.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (R + k >= n_own) len[k] = 0;
          if (k < 3 && k == k0 && acc0) len[k] = c0[u];
          if (k == 3 && acc1) len[k] = c1[u];
          sum += len[k];
        }
      }
      s_len[u*RB_THREADS + t] = sum;
    }
    __syncthreads();

    // ---- per-branch tile scan + chained look-back: one warp per branch of the chunk (BC <= 8) --------------
    constexpr int PER = RB_THREADS/32;
    if (warp < nb) {
      const int bi = b0 + warp;
      uint32_t *row = s_len + warp*RB_THREADS;
      uint32_t v[PER], mine = 0;
#pragma unroll
      for (int k = 0; k < PER; ++k) { v[k] = row[lane*PER + k]; mine += v[k]; }
      uint32_t incl = mine;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += up;
      }
      const uint32_t tot = __shfl_sync(0xffffffffu, incl, 31);
      const uint32_t base = tile_base(a.desc + (uint64_t)bi*a.n_tiles, tile, a.epoch, tot, lane);
      if (lane == 0 && (uint64_t)base + tot > a.jcap[bi]) {
        // nothing more of this tile is written; the scans go on (later tiles wait for every branch's
        // descriptor), so the last tile records the branch's new total and the host can regrow and rerun
        atomicOr(a.err, ERR_JUMP_CAPACITY);
        atomicMax(a.jneed + bi, (unsigned long long)base + tot);
        s_fail = 1;
      }
      uint32_t run = base + incl - mine;
#pragma unroll
      for (int k = 0; k < PER; ++k) { row[lane*PER + k] = run; run += v[k]; }
    }
    __syncthreads();
    const bool write = !s_fail;

    // ---- the rebuilt rows: jump times to HBM, offsets IN PLACE in shared memory ---------------------------------
    // A thread owns elements 4t..4t+3 of the staged run; the fifth input (the next quad's first offset) comes
    // from the neighbouring lane's registers, or from s_edge across a warp boundary, so no thread reads an
    // element another thread overwrites.
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      if (u >= nb) break;
      const int bi = b0 + u;
      uint32_t *o = so_base + (size_t)u*ROW_ELEMS + 4*t;
      const uint4 q = *reinterpret_cast<const uint4 *>(o);
      uint32_t e = __shfl_down_sync(0xffffffffu, q.x, 1);
      if (lane == 31) e = s_edge[u*(RB_THREADS/32) + warp];
      if (!write || R > n_own) continue;
      uint32_t w = s_len[u*RB_THREADS + t];
      uint32_t out[4] = {w, w, w, w};                  // R == n_own: the closing offset
      if (R < n_own) {
        const uint32_t in[5] = {q.x, q.y, q.z, q.w, e};
        const uint64_t jb = a.in.jbase[bi];
        const double *src_base = a.in.jmp + jb;
        double *dst_base = a.jmp_out + jb;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          out[k] = w;
          if (R + k >= n_own) continue;
          if (k < 3 && k == k0 && acc0) {
            const int n_new = (int)c0[u];
            for (int x = 0; x < n_new; ++x) dst_base[w + x] = pj0[poff0 + x];
            poff0 += n_new; w += n_new;
          }
          else if (k == 3 && acc1) {
            const int n_new = (int)c1[u];
            for (int x = 0; x < n_new; ++x) dst_base[w + x] = pj1[poff1 + x];
            poff1 += n_new; w += n_new;
          }
          else {
            const uint32_t n = in[k + 1] - in[k];
            for (uint32_t x = 0; x < n; ++x) dst_base[w + x] = src_base[in[k] + x];
            w += n;
          }
        }
      }
      *reinterpret_cast<uint4 *>(o) = make_uint4(out[0], out[1], out[2], out[3]);
    }
    fence_async_smem();               // the generic-proxy writes above, before the copy engine reads them
    __syncthreads();
    if (t == 0 && write) {
      for (int u = 0; u < nb; ++u)
        store_async(a.off_out + (uint64_t)(b0 + u)*a.in.off_pitch + R0, so_base + (size_t)u*ROW_ELEMS, store_elems*4u);
      store_commit();
    }
  }
  if (t == 0) store_wait_read_all();   // shared memory must outlive the copy engine's reads
}

}  // namespace

size_t
mh_pass_smem_bytes(int n_nodes) {
  return (size_t)(n_nodes - 1)*RB_THREADS*sizeof(uint32_t);
}

// A sharded context reads its neighbours' boundary rows from the halo store; only the few
// groups within two sites of the owned block's ends can touch it.  Those run in a small
// launch of the halo-aware instantiation (launch_mh_edge), everything else in the halo-free one
// (launch_mh_interior; 15 % faster: 5.5 vs 6.5 ms per pass).  The two launches touch disjoint groups
// and may run concurrently on two streams (the multi-GPU driver starts the edge groups first so
// that their rows can travel to the neighbour GPU while the interior is still being updated).
static void
edge_split(const PassArgs &a, uint64_t &g_lo, uint64_t &g_hi) {
  const uint64_t lo = a.in.own_lo, hi = a.in.own_hi;
  g_lo = lo == 0 ? 0 : (lo + 4)/3;                                 // first group with 3g - 2 >= own_lo
  g_hi = hi == a.in.n_local ? a.n_groups : (hi >= 5 ? (hi - 5)/3 + 1 : 0);   // 3g + 4 < own_hi
  g_lo = g_lo < a.n_groups ? g_lo : a.n_groups;
  g_hi = g_hi > g_lo ? g_hi : g_lo;
  g_hi = g_hi < a.n_groups ? g_hi : a.n_groups;
}

// Shared memory per CTA that still lets mh_minblocks(MAXN, HALO) CTAs share an SM (227 KB usable, 1 KB
// reserved per CTA, ~0.4 KB static).
template <int MAXN, bool HALO>
static size_t smem_budget() { return (size_t)(232448 - 1024*mh_minblocks(MAXN, HALO))/mh_minblocks(MAXN, HALO) - 512; }

template <int MAXN, int MAXJ, bool HALO, bool DEFER>
static void
launch_variant(const PassArgs &a, unsigned blocks, cudaStream_t st) {
  constexpr int MT = mh_threads(MAXN, HALO);
  size_t off[7];
  const bool qsmem = EPV_MH_QSMEM && mh_lock(MAXN) && mh_smem_layout<MAXN, MT>(a.tree.n_nodes, a.q_slots, off) <= smem_budget<MAXN, HALO>();
  if (qsmem) {
    const size_t smem = mh_smem_layout<MAXN, MT>(a.tree.n_nodes, a.q_slots, off);
    cudaFuncSetAttribute(mh_update_kernel<MAXN, MAXJ, HALO, DEFER, mh_lock(MAXN)>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    mh_update_kernel<MAXN, MAXJ, HALO, DEFER, mh_lock(MAXN)><<<blocks, MT, smem, st>>>(a, a.q_slots);
  }
  else {
    const size_t smem = mh_smem_layout<MAXN, MT>(a.tree.n_nodes, 0, off);
    cudaFuncSetAttribute(mh_update_kernel<MAXN, MAXJ, HALO, DEFER, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    mh_update_kernel<MAXN, MAXJ, HALO, DEFER, false><<<blocks, MT, smem, st>>>(a, 0);
  }
}

template <int MAXN, int MAXJ>
static int
launch_interior_t(const PassArgs &a0, cudaStream_t st) {
  PassArgs a = a0;
  a.list_mode = 0;
  a.reverse = getenv("EPIEVO_B200_REVERSE") != nullptr;
  uint64_t g_lo, g_hi; edge_split(a, g_lo, g_hi);
  if (g_hi <= g_lo) return 0;
  // deferral costs one more launch whose length is the latency of a single site update (~0.1 ms
  // when a handful of sites overflow): worth it only when the pass is long (1e6-site single branch:
  // 0.45 vs 0.41 ms per pass with it)
  // (EPIEVO_B200_DEFER_MIN overrides the threshold: the tests run both ways on small inputs)
  const char *env = getenv("EPIEVO_B200_DEFER_MIN");
  const uint64_t defer_min = env ? strtoull(env, nullptr, 10) : (1ull << 20);
  const bool defer = mh_defer(MAXN) && g_hi - g_lo >= defer_min;
  a.g_first = g_lo; a.g_count = a.g_split = g_hi - g_lo; a.g_skip = 0;
  const unsigned blocks = (unsigned)((a.g_count + mh_threads(MAXN, false) - 1)/mh_threads(MAXN, false));
  if (defer) launch_variant<MAXN, MAXJ, false, mh_defer(MAXN)>(a, blocks, st);
  else launch_variant<MAXN, MAXJ, false, false>(a, blocks, st);
  if (!defer) return 1;
  // the sites the main launch deferred (list length known on the device only)
  a.list_mode = 1;
  launch_variant<MAXN, MAXJ, true, false>(a, MH_OVF_BLOCKS, st);
  return 2;
}

template <int MAXN, int MAXJ>
static int
launch_edge_t(const PassArgs &a0, cudaStream_t st) {
  PassArgs a = a0;
  a.list_mode = 0;
  uint64_t g_lo, g_hi; edge_split(a, g_lo, g_hi);
  const uint64_t n_edge = g_lo + (a.n_groups - g_hi);
  if (!n_edge) return 0;
  a.g_first = 0; a.g_split = g_lo; a.g_skip = g_hi - g_lo; a.g_count = n_edge;
  launch_variant<MAXN, MAXJ, true, false>(a, (unsigned)((n_edge + mh_threads(MAXN, true) - 1)/mh_threads(MAXN, true)), st);
  return 1;
}

int
launch_mh_interior(const PassArgs &a, cudaStream_t st) {
  const int n = a.tree.n_nodes;
  if (n <= 8) return launch_interior_t<8, 16>(a, st);
  if (n <= 16) return launch_interior_t<16, 32>(a, st);
  if (n <= 64) return launch_interior_t<64, 64>(a, st);
  return launch_interior_t<256, 128>(a, st);
}

int
launch_mh_edge(const PassArgs &a, cudaStream_t st) {
  const int n = a.tree.n_nodes;
  if (n <= 8) return launch_edge_t<8, 16>(a, st);
  if (n <= 16) return launch_edge_t<16, 32>(a, st);
  if (n <= 64) return launch_edge_t<64, 64>(a, st);
  return launch_edge_t<256, 128>(a, st);
}

int
launch_mh_pass(const PassArgs &a, cudaStream_t st) {
  const int n = launch_mh_interior(a, st);
  return n + launch_mh_edge(a, st);
}

// EPIEVO_B200_REBUILD=bulk[:BC] selects the bulk-staged kernel (BC = branches per chunk, 1..8); the default is
// the kernel with per-thread loads.  (Read once; an experiment switch, not an API.)
static int
rebuild_bulk_chunk(int n_br) {
  static int mode = -2;
  if (mode == -2) {
    const char *e = getenv("EPIEVO_B200_REBUILD");
    mode = -1;
    if (e && !strncmp(e, "bulk", 4)) mode = e[4] == ':' ? atoi(e + 5) : 0;
  }
  if (mode < 0) return 0;
  int bc = mode > 0 ? mode : 4;
  bc = bc > 8 ? 8 : bc;
  return bc > n_br ? n_br : bc;
}

void
launch_csr_rebuild(const PassArgs &a, cudaStream_t st) {
  const int n_br = a.tree.n_nodes - 1;
  if (const int bc = rebuild_bulk_chunk(n_br)) {
    const size_t smem = ((size_t)2*bc*bulk::ROW_ELEMS + (size_t)bc*RB_THREADS + (size_t)bc*(RB_THREADS/32))*sizeof(uint32_t)
                      + 2*sizeof(uint64_t);
    cudaFuncSetAttribute(csr_rebuild_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    csr_rebuild_bulk_kernel<<<a.n_tiles, RB_THREADS, smem, st>>>(a, bc);
    return;
  }
  const size_t smem = mh_pass_smem_bytes(a.tree.n_nodes);
  // branches whose loads are grouped: measured best 2 for 14 branches (1.03 vs 1.07 / 1.13 ms at
  // 1 / 4), 4 for 62 (1.52 vs 1.67 / 1.60 ms at 1 / 2)
  if (a.tree.n_nodes <= 16) {
    cudaFuncSetAttribute(csr_rebuild_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    csr_rebuild_kernel<2><<<a.n_tiles, RB_THREADS, smem, st>>>(a);
  }
  else {
    cudaFuncSetAttribute(csr_rebuild_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    csr_rebuild_kernel<4><<<a.n_tiles, RB_THREADS, smem, st>>>(a);
  }
}

}  // namespace epv
