// kernels_stats.cu -- J/D triplet sufficient statistics, log-likelihood / gradient,
// jump-time rescale and the host-format import/export kernels.
#include "kernels.h"
#include "rows.cuh"

namespace epv {

namespace {

constexpr int SS_THREADS = 256;

// Per-branch J/D over every interior site of the shard
// (get_sufficient_statistics, ParamEstimation.cpp:92-114, built on
// add_sufficient_statistics, Path.cpp:267-301).
//
// lane = site; a warp walks chunks of 32 x SS_ITERS consecutive sites, SS_GROUP branches
// at a time.  Two kinds of rows:
//  * event-free rows (the three paths have no jump on the branch; 85-98 %) only add
//    tot_time to D[triplet]: each lane counts them in eight 8-bit fields of one 64-bit
//    register per branch, reduced once per (chunk, branch) with a warp shuffle tree;
//  * rows with events are NOT merged inline (that would run the three-way merge with a
//    few lanes active on almost every iteration): the lane appends (site, branch) to a
//    per-warp shared-memory queue (ballot + popc), and the warp merges queued rows 32 at
//    a time, every lane one row (same strict-'<' tie-breaks as the reference).  Their D
//    terms are accumulated in shared memory as 64-bit fixed point (quantum
//    tot_time * 2^-40: native integer atomics instead of fp64 CAS loops; the rounding is
//    <= 2^-41 of a branch length per term, far inside the 1e-9 parity bound), J as
//    integers.
// One global atomic per (branch, triplet) per CTA, all of them integer, so the result does not
// depend on the order in which CTAs (or the tile / generic kernels) finish.  acc layout per
// branch: [0..8) low 32 bits of the fixed-point D sums, [8..16) J counts, [16..24) event-free
// row counts, [24..32) high bits of the D sums, [32..56) / [56..80) low / high bits of the tile
// kernel's event-time sums by class (see there).
constexpr int SS_ITERS = 16;     // warp rows (of 32 sites) per chunk
constexpr int SS_GROUP = 4;   // measured 1 / 2 / 3 / 4 / 8 / 14: 1.97 / 1.51 / 1.36 / 1.31 / 1.42 / 1.76 ms (load parallelism vs code size)      // branches whose packed counters live in registers at once
constexpr int SS_QCAP = 96;      // queue entries per warp (drained in batches of 32)
constexpr double SS_FIX = 1099511627776.0;   // 2^40
constexpr int SS_ACC = 80;       // accumulator words per branch

struct StatsAcc {
  long long *Dfix;     // [n_br*8] fixed-point D of rows with events
  unsigned *J;         // [n_br*8]
  unsigned *C;         // [n_br*8] event-free rows
};

__device__ __forceinline__ void
flush_acc(const StatsAcc &A, int i, unsigned long long *acc) {
  const int bi = i >> 3, t = i & 7;
  const long long d = A.Dfix[i];                 // signed: an event's time is added to one bin, taken from another
  if (d) {
    atomicAdd(acc + bi*SS_ACC + t, (unsigned long long)d & 0xffffffffull);
    atomicAdd(acc + bi*SS_ACC + 24 + t, (unsigned long long)(d >> 32));
  }
  if (A.J[i]) atomicAdd(acc + bi*SS_ACC + 8 + t, (unsigned long long)A.J[i]);
  if (A.C[i]) atomicAdd(acc + bi*SS_ACC + 16 + t, (unsigned long long)A.C[i]);
}

__device__ __forceinline__ void
load3(const PathView &pv, const uint8_t *state, int W, int bi, uint64_t s, Row &L, Row &M, Row &R) {
  if (s - 1 >= pv.own_lo && s + 1 < pv.own_hi && W <= 8) {
    const uint32_t *o = pv.off + (uint64_t)bi*pv.off_pitch + (s - 1 - pv.own_lo);
    const uint32_t o0 = o[0], o1 = o[1], o2 = o[2], o3 = o[3];
    const double *jb = pv.jmp + pv.jbase[bi];
    L.n = (int)(o1 - o0); M.n = (int)(o2 - o1); R.n = (int)(o3 - o2);
    L.ptr = jb + o0; M.ptr = jb + o1; R.ptr = jb + o2;
    L.init = (int)((state_word(state, W, s - 1) >> (bi + 1)) & 1);
    M.init = (int)((state_word(state, W, s) >> (bi + 1)) & 1);
    R.init = (int)((state_word(state, W, s + 1) >> (bi + 1)) & 1);
  }
  else {
    L = load_row(pv, state, W, bi, s - 1);
    M = load_row(pv, state, W, bi, s);
    R = load_row(pv, state, W, bi, s + 1);
  }
}

// Three-way merge of one row triple (add_sufficient_statistics, Path.cpp:267-301).  D is accumulated
// through the same identity and the SAME integers as the tile kernel below -- an event at time t adds
// fix(t) to the bin of the triplet before it and takes it from the bin of the triplet after it, the final
// triplet goes to the row counter C (the fold multiplies it by the branch length) -- so the integer sums,
// hence J and D, do not depend on which kernel a row went to (i.e. on where the shard ends are).
__device__ __forceinline__ void
merge_row(const Row &L, const Row &M, const Row &R, double tot, const StatsAcc &A, int bi) {
  int trip = L.init*4 + M.init*2 + R.init, i = 0, j = 0, k = 0;
  const double scale = SS_FIX/tot;
  const double INF = __longlong_as_double(0x7ff0000000000000LL);
  for (;;) {
    const double tl = i < L.n ? L.ptr[i] : INF;
    const double tm = j < M.n ? M.ptr[j] : INF;
    const double tr = k < R.n ? R.ptr[k] : INF;
    double tn; int flip;
    if (tl < fmin(tm, tr)) { tn = tl; flip = 4; ++i; }
    else if (tm < tr) { tn = tm; flip = 2; ++j; }
    else { tn = tr; flip = 1; ++k; }
    if (!(tn < INF)) { atomicAdd(&A.C[bi*8 + trip], 1u); break; }
    const unsigned long long f = (unsigned long long)__double2ll_rn(tn*scale);
    atomicAdd(reinterpret_cast<unsigned long long *>(&A.Dfix[bi*8 + trip]), f);
    atomicAdd(reinterpret_cast<unsigned long long *>(&A.Dfix[bi*8 + (trip ^ flip)]), 0ull - f);
    if (flip == 2) atomicAdd(&A.J[bi*8 + trip], 1u);
    trip ^= flip;
  }
}

static __device__ __noinline__ void
ss_drain(const PathView &pv, const uint8_t *state, int W, const double *branch, const StatsAcc &A,
         const unsigned *queue, int n_take, uint64_t c0) {
  const int lane = threadIdx.x & 31;
  __syncwarp();
  if (lane < n_take) {
    const unsigned it = queue[lane];
    const int bi = (int)(it >> 16);
    const uint64_t s = c0 + (it & 0xffffu);
    Row L, M, R;
    load3(pv, state, W, bi, s, L, M, R);
    merge_row(L, M, R, branch[bi + 1], A, bi);
  }
  __syncwarp();
}

__global__ void __launch_bounds__(SS_THREADS, 3)
suffstats_kernel(const PathView pv, const uint8_t *state, int W, int n_br, const double *branch,
                 uint64_t lo, uint64_t hi, unsigned long long *acc) {
  extern __shared__ unsigned long long s_raw[];
  StatsAcc A;
  A.Dfix = reinterpret_cast<long long *>(s_raw);                  // [n_br*8]
  A.J = reinterpret_cast<unsigned *>(s_raw + n_br*8);             // [n_br*8]
  A.C = A.J + n_br*8;                                             // [n_br*8]
  unsigned *s_queue = A.C + n_br*8 + (threadIdx.x >> 5)*SS_QCAP;  // this warp's queue
  for (int i = threadIdx.x; i < n_br*8; i += SS_THREADS) { A.Dfix[i] = 0; A.J[i] = 0; A.C[i] = 0; }
  __syncthreads();
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const uint64_t warp_id = ((uint64_t)blockIdx.x*SS_THREADS + threadIdx.x) >> 5;
  const uint64_t n_warps = ((uint64_t)gridDim.x*SS_THREADS) >> 5;
  const uint64_t chunk = 32*SS_ITERS;
  const uint64_t n_chunks = (hi - lo + chunk - 1)/chunk;
  for (uint64_t ch = warp_id; ch < n_chunks; ch += n_warps) {
    const uint64_t c0 = lo + ch*chunk;
    int qn = 0;                                  // queue fill (warp-uniform)
    // merge 32 queued rows (or the last qn), one per lane -- out of line: the merge is large
    // and inlining it at every call site made the kernel 104 KB of SASS, with instruction
    // fetch as its first stall reason
    auto drain = [&](int n_take) {
      ss_drain(pv, state, W, branch, A, s_queue + (qn - n_take), n_take, c0);
      qn -= n_take;
    };
    for (int b0 = 0; b0 < n_br; b0 += SS_GROUP) {
      uint64_t packed[SS_GROUP];                 // per branch: eight 8-bit counters of event-free rows
#pragma unroll
      for (int g = 0; g < SS_GROUP; ++g) packed[g] = 0;
      for (int it = 0; it < SS_ITERS; ++it) {
        if (c0 + 32*it >= hi) break;             // warp-uniform
        const uint64_t s = c0 + 32*it + lane;
        const bool live = s < hi;
        const bool fast = live && s - 1 >= pv.own_lo && s + 1 < pv.own_hi && W <= 8;
        uint64_t wl = 0, wm = 0, wr = 0;
        if (fast) { wl = state_word(state, W, s - 1); wm = state_word(state, W, s); wr = state_word(state, W, s + 1); }
        // issue the offset loads of the whole branch group first (memory-level parallelism:
        // the pass is latency-bound, not bandwidth-bound), then classify
        uint32_t o0[SS_GROUP], o3[SS_GROUP];
#pragma unroll
        for (int g = 0; g < SS_GROUP; ++g) {
          o0[g] = 0; o3[g] = 1;
          if (fast && b0 + g < n_br) {
            const uint32_t *o = pv.off + (uint64_t)(b0 + g)*pv.off_pitch + (s - 1 - pv.own_lo);
            o0[g] = o[0]; o3[g] = o[3];
          }
        }
#pragma unroll
        for (int g = 0; g < SS_GROUP; ++g) {
          const int bi = b0 + g;
          if (bi >= n_br) break;
          bool slow = false;
          if (fast) {
            if (o0[g] == o3[g])
              packed[g] += 1ull << (8*((int)((wl >> (bi + 1)) & 1)*4 + (int)((wm >> (bi + 1)) & 1)*2 +
                                       (int)((wr >> (bi + 1)) & 1)));
            else slow = true;
          }
          else slow = live;                      // halo neighbourhood / wide tree: generic loads
          const unsigned m = __ballot_sync(FULL, slow);
          if (m) {
            if (slow) s_queue[qn + __popc(m & ((1u << lane) - 1u))] = ((unsigned)bi << 16) | (unsigned)(32*it + lane);
            qn += __popc(m);
            if (qn > SS_QCAP - 32) drain(32);
          }
        }
      }
      // reduce the packed counters over the warp: a field is <= 16 per lane, so two words
      // of four 16-bit fields hold the sums over 32 lanes
#pragma unroll
      for (int g = 0; g < SS_GROUP; ++g) {
        const int bi = b0 + g;
        if (bi >= n_br) break;
        uint64_t ev = packed[g] & 0x00ff00ff00ff00ffull, od = (packed[g] >> 8) & 0x00ff00ff00ff00ffull;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
          ev += __shfl_xor_sync(FULL, ev, d);
          od += __shfl_xor_sync(FULL, od, d);
        }
        if (lane < 8) {
          const uint64_t src = (lane & 1) ? od : ev;
          const unsigned c = (unsigned)((src >> (16*(lane >> 1))) & 0xffffu);
          if (c) atomicAdd(&A.C[bi*8 + lane], c);
        }
      }
    }
    while (qn > 0) drain(qn < 32 ? qn : 32);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n_br*8; i += SS_THREADS) {
    flush_acc(A, i, acc);
  }
}

// ---- tile kernel: the interior of the shard, 128 rows x one branch per step ----------------
// Works from the identity (segment lengths telescope over the events of a row triple)
//   D[trip] = tot * #{triples whose FINAL triplet is trip}
//             + sum over events at time t of t * ([triplet before = trip] - [triplet after = trip]),
// so no row triple is ever merged:
//  * every row is classified by its final triplet with masks and popcounts.  A warp owns a tile
//    of 128 consecutive rows; lane i holds rows 4i..4i+3 (one 16-byte load of the offsets per
//    branch).  Bit i of mask j = row 4i+j ("strided" order: the right neighbour of (i, j<3) is
//    (i, j+1), of (i, 3) it is (i+1, 0), so a neighbour mask is another mask or a one-bit shift
//    of one, never a bit permutation).  F_j = ballot(initial state bit) ^ ballot(jump count odd)
//    is the final state; lane k < 8 counts the rows of triplet k with four popcounts (pattern
//    match = three xors against per-lane constants);
//  * only rows that HOLD jumps (about 5 %; a row triple with an event is about 14 %) go to the
//    per-warp queue.  A queued row is taken by one lane, which visits its jumps: the jump at time
//    t is an event of the three triples centred on sites s-1, s, s+1; their triplets just before
//    t come from the initial states of s-2..s+2 and, for the few neighbours that hold jumps
//    themselves, the parity of their jumps before t (ties: the higher site first, as
//    Path.cpp:275-281).  One fixed-point add of t per centre into the bin of (triplet before,
//    which row jumped); suffstats_fold_kernel applies the +/-.  Bins are 64-bit integers in
//    shared memory (CAS loops on this architecture), replicated in columns picked by the lane so
//    that the lanes of a warp rarely meet in one bin.
// The rows next to the shard's ends (and trees wider than 64 nodes) stay with suffstats_kernel.
constexpr int ST_THREADS = 256;
constexpr int ST_ROWS = 128;     // rows per tile
constexpr int ST_QCAP = 168;     // < 32 left over + up to 130 new entries per (tile, branch)
#ifndef ST_MINB_
#define ST_MINB_ 4
#endif
#ifndef ST_BINKB_
#define ST_BINKB_ 24
#endif
#ifndef ST_WAVES_
#define ST_WAVES_ 2
#endif
#ifndef ST_PIPE_
#define ST_PIPE_ 0
#endif
__device__ __forceinline__ void prefetch_l1(const void *p) { asm volatile("prefetch.global.L1 [%0];" :: "l"(p)); }

// ST_SPLIT_ (default): the 40-bit fixed-point event times go to the bins as three 14-bit pieces with native
// 32-bit shared atomics (RED, fire and forget); a 64-bit shared atomicAdd is a compare-and-swap spin loop on
// this architecture (ATOMS.CAST.SPIN.64) and was a third of the kernel's instructions.  A piece bin holds
// 2^18 adds; every lane counts its adds and after 1024 (256 lanes x 1024 = 2^18) goes on in the 64-bit
// bins, so the sums stay exact whatever the data.
#ifndef ST_SPLIT_
#define ST_SPLIT_ 1
#endif
constexpr unsigned ST_LANE_ADDS = 1024;

struct TileShared {
  long long *S;        // ST_SPLIT_: [n_br*24] 64-bit overflow bins; else [n_br*24*cols] fixed-point sums of event
                       // times, class = trip_before*3 + which
  unsigned *P;         // ST_SPLIT_: [n_br*24*cols][3] sums of the 14-bit pieces
  unsigned *J;         // [n_br*8]
  unsigned *C;         // [n_br*8] rows by final triplet
  double *scale;       // [n_br] SS_FIX / tot_time
  int col_shift;       // log2(cols)
  uint64_t c_lo, c_hi; // sites whose triple this kernel accounts for
};

// Offsets of one (tile, branch) step: lane i holds rows 4i..4i+3; lane 0 also the offset before the
// tile, lane 31 the two after it.
struct TileStep { uint4 o; uint32_t el; uint2 er; };

__device__ __forceinline__ TileStep
load_step(const uint32_t *ob, int lane) {      // ob = the branch's offset of the tile's first row
  TileStep r;
  r.o = *reinterpret_cast<const uint4 *>(ob + 4*lane);
  r.el = 0; r.er = make_uint2(0, 0);
  if (lane == 0) r.el = *(ob - 1);
  if (lane == 31) r.er = *reinterpret_cast<const uint2 *>(ob + ST_ROWS);
  return r;
}

// One event time into the bin of (branch, class).
__device__ __forceinline__ void
bin_add(const TileShared &T, int bi, int cls, int col, unsigned long long v, unsigned &n_adds) {
#if ST_SPLIT_
  if (n_adds < ST_LANE_ADDS) {
    unsigned *p = T.P + 3*((((size_t)bi*24 + cls) << T.col_shift) + col);
    atomicAdd(p, (unsigned)(v & 0x3fffu));
    atomicAdd(p + 1, (unsigned)((v >> 14) & 0x3fffu));
    atomicAdd(p + 2, (unsigned)(v >> 28));
    ++n_adds;
  }
  else atomicAdd(reinterpret_cast<unsigned long long *>(T.S + bi*24 + cls), v);
#else
  atomicAdd(reinterpret_cast<unsigned long long *>(T.S + ((((size_t)bi*24 + cls) << T.col_shift) + col)), v);
#endif
}

// (one call site, inlined: as an out-of-line call its by-reference arguments and the registers saved
// around it were local-memory traffic twice the size of the kernel's global loads)
__device__ __forceinline__ void
st_drain(const PathView &pv, const uint8_t *state, int W, const TileShared &T, const unsigned *queue,
         int n_take, uint64_t first_s0, uint64_t ord_stride, unsigned &n_adds) {
  const int lane = threadIdx.x & 31;
  __syncwarp();
  if (lane < n_take) {
    const unsigned it = queue[lane];
    const int bi = (int)(it >> 26);
    // code 0 = the row before the tile, 1..128 = its rows, 129 = the row after it
    const uint64_t s = first_s0 + (uint64_t)((it >> 8) & 0x3ffffu)*ord_stride + (it & 0xffu) - 1;
    const bool vL = s - 1 >= T.c_lo && s - 1 < T.c_hi, vM = s >= T.c_lo && s < T.c_hi,
               vR = s + 1 >= T.c_lo && s + 1 < T.c_hi;
    const uint32_t *o = pv.off + (uint64_t)bi*pv.off_pitch + (s - pv.own_lo);
    const uint32_t o0 = o[0], o1 = o[1];
    // rows s-2, s-1, s, s+1, s+2 are [a2, a1), [a1, o0), [o0, o1), [o1, c2), [c2, e2) of the branch's
    // jump times; a neighbour that no valid centre needs gets an empty range
    uint32_t a1 = o0, a2 = o0, c2 = o1, e2 = o1;
    int b2 = 0, b1 = 0, d1 = 0, d2 = 0;       // initial states of s-2, s-1, s+1, s+2
    const int sh = bi + 1;
    if (vL || vM) { a1 = o[-1]; a2 = a1; b1 = (int)((state_word(state, W, s - 1) >> sh) & 1); }
    if (vL) { a2 = o[-2]; b2 = (int)((state_word(state, W, s - 2) >> sh) & 1); }
    if (vM || vR) { c2 = o[2]; e2 = c2; d1 = (int)((state_word(state, W, s + 1) >> sh) & 1); }
    if (vR) { e2 = o[3]; d2 = (int)((state_word(state, W, s + 2) >> sh) & 1); }
    const int b0 = (int)((state_word(state, W, s) >> sh) & 1);
    const double *jb = pv.jmp + pv.jbase[bi];
    const double scale = T.scale[bi];
    const int col = lane & ((1 << T.col_shift) - 1);
    const uint32_t nl = o0 - a2, nn = nl + (e2 - o1);
    for (uint32_t k = 0; k < o1 - o0; ++k) {
      const double t = jb[o0 + k];
      const unsigned long long v = (unsigned long long)__double2ll_rn(t*scale);
      const int own = b0 ^ (int)(k & 1);
      // parity of the neighbours' jumps before t, one flat loop over all four rows (ties: a row to
      // the right jumps first)
      unsigned fl = 0;
      for (uint32_t j = 0; j < nn; ++j) {
        const bool left = j < nl;
        const uint32_t i = left ? a2 + j : o1 + (j - nl);
        const double q = jb[i];
        const bool before = left ? q < t : q <= t;
        fl ^= (unsigned)before << (left ? (int)(i >= a1) : 2 + (int)(i >= c2));
      }
      const int x2 = b2 ^ (int)(fl & 1), x1 = b1 ^ (int)((fl >> 1) & 1), y1 = d1 ^ (int)((fl >> 2) & 1),
                y2 = d2 ^ (int)(fl >> 3);
      if (vL) bin_add(T, bi, (x2*4 + x1*2 + own)*3 + 0, col, v, n_adds);
      if (vM) {
        const int tb = x1*4 + own*2 + y1;
        bin_add(T, bi, tb*3 + 1, col, v, n_adds);
        atomicAdd(&T.J[bi*8 + tb], 1u);
      }
      if (vR) bin_add(T, bi, (own*4 + y1*2 + y2)*3 + 2, col, v, n_adds);
    }
  }
  __syncwarp();
}

template <typename WT>           // state bits of one site: uint32_t up to 32 nodes, else uint64_t
__global__ void __launch_bounds__(ST_THREADS, ST_MINB_)
suffstats_tile_kernel(const PathView pv, const uint8_t *state, int W, int n_br, const double *branch,
                      uint64_t r0, uint64_t n_tiles, int col_shift, unsigned long long *acc) {
  extern __shared__ unsigned long long s_raw[];
  TileShared T;
#if ST_SPLIT_
  // [n_br*24] u64 overflow bins, [n_br] scales, then the piece bins (u32 x 3 per bin and column), J, C, queues
  const int n_bins = n_br*24, n_pieces = 3*((n_br*24) << col_shift);
  T.S = reinterpret_cast<long long *>(s_raw);
  T.scale = reinterpret_cast<double *>(s_raw + n_bins);
  T.P = reinterpret_cast<unsigned *>(s_raw + n_bins + n_br);
  T.J = T.P + n_pieces;
  for (int i = threadIdx.x; i < n_pieces; i += ST_THREADS) T.P[i] = 0;
#else
  const int n_bins = (n_br*24) << col_shift;
  T.S = reinterpret_cast<long long *>(s_raw);
  T.scale = reinterpret_cast<double *>(s_raw + n_bins);
  T.J = reinterpret_cast<unsigned *>(s_raw + n_bins + n_br);
  T.P = nullptr;
#endif
  T.C = T.J + n_br*8;
  T.col_shift = col_shift;
  T.c_lo = pv.own_lo + r0; T.c_hi = T.c_lo + n_tiles*ST_ROWS;
  unsigned *s_queue = T.C + n_br*8 + (threadIdx.x >> 5)*ST_QCAP;
  for (int i = threadIdx.x; i < n_bins; i += ST_THREADS) T.S[i] = 0;
  for (int i = threadIdx.x; i < n_br*8; i += ST_THREADS) { T.J[i] = 0; T.C[i] = 0; }
  for (int i = threadIdx.x; i < n_br; i += ST_THREADS) T.scale[i] = SS_FIX/branch[i + 1];
  __syncthreads();
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u;
  const int k = lane & 7;
  const unsigned xl = (k & 4) ? 0u : ~0u, xm = (k & 2) ? 0u : ~0u, xr = (k & 1) ? 0u : ~0u;
  const uint64_t warp_id = ((uint64_t)blockIdx.x*ST_THREADS + threadIdx.x) >> 5;
  const uint64_t n_warps = ((uint64_t)gridDim.x*ST_THREADS) >> 5;
  // tile = warp_id + ord * n_warps: neighbouring tiles are read by neighbouring warps at about the
  // same time (a contiguous span per warp was tried: 66 000 separate 512-byte streams, DRAM read
  // traffic and latency both up).  Queue entries carry ord, so the queue runs on across tiles and
  // only the last batch of a warp is partial.
  const uint64_t first_s0 = pv.own_lo + r0 + warp_id*ST_ROWS, ord_stride = n_warps*ST_ROWS;
  int qn = 0;                                    // queue fill (warp-uniform)
  unsigned n_adds = 0;                           // this lane's adds to the piece bins (bin_add)
  TileStep cur;
  const uint32_t *p_tile = pv.off + r0 + warp_id*ST_ROWS;    // branch 0, first row of the warp's tile
#if ST_PIPE_
  if (warp_id < n_tiles) cur = load_step(p_tile, lane);
#endif
  unsigned ord = 0;
  for (uint64_t tile = warp_id; tile < n_tiles; tile += n_warps, ++ord, p_tile += ord_stride) {
    const uint32_t *p_step = p_tile;
    const uint64_t t0 = r0 + tile*ST_ROWS;       // first row of the tile (multiple of 4, >= 4)
    const uint64_t s0 = pv.own_lo + t0;          // its site
    const WT w0 = (WT)state_word(state, W, s0 + 4*lane), w1 = (WT)state_word(state, W, s0 + 4*lane + 1),
             w2 = (WT)state_word(state, W, s0 + 4*lane + 2), w3 = (WT)state_word(state, W, s0 + 4*lane + 3);
    const WT we_l = (WT)state_word(state, W, s0 - 1), we_r = (WT)state_word(state, W, s0 + ST_ROWS);
    const unsigned tcode = ord << 8;
    // the rows just outside the kernel's range hold events of the first / last triple inside it
    const bool edge_l = tile == 0, edge_r = tile + 1 == n_tiles;
    for (int bi = 0; bi < n_br; ++bi) {
#if ST_PIPE_
      // the loads of the next step (next branch, or the first branch of the warp's next tile) are
      // issued before this step's are used
      const uint4 o = cur.o; const uint32_t el = cur.el; const uint2 er = cur.er;
      {
        const bool wrap = bi + 1 == n_br;
        p_step = wrap ? p_tile + ord_stride : p_step + pv.off_pitch;
        if (!wrap || tile + n_warps < n_tiles) cur = load_step(p_step, lane);
      }
#else
      cur = load_step(p_step, lane);
      p_step += pv.off_pitch;
      const uint4 o = cur.o; const uint32_t el = cur.el; const uint2 er = cur.er;
#endif
      uint32_t o4 = __shfl_down_sync(FULL, o.x, 1);
      if (lane == 31) o4 = er.x;
      const uint32_t n0 = o.y - o.x, n1 = o.z - o.y, n2 = o.w - o.z, n3 = o4 - o.w;
      const unsigned J0 = __ballot_sync(FULL, n0 != 0), J1 = __ballot_sync(FULL, n1 != 0),
                     J2 = __ballot_sync(FULL, n2 != 0), J3 = __ballot_sync(FULL, n3 != 0);
      // jump counts of rows t0-1 (lane 0) and t0+128 (lane 31)
      const uint32_t ne = lane == 0 ? o.x - el : (lane == 31 ? er.y - er.x : 0u);
      const unsigned pe = __ballot_sync(FULL, ne & 1);
      const int sh = bi + 1;
      unsigned F0 = __ballot_sync(FULL, ((w0 >> sh) ^ n0) & 1), F1 = __ballot_sync(FULL, ((w1 >> sh) ^ n1) & 1),
               F2 = __ballot_sync(FULL, ((w2 >> sh) ^ n2) & 1), F3 = __ballot_sync(FULL, ((w3 >> sh) ^ n3) & 1);
      const unsigned L0 = (F3 << 1) | (((unsigned)(we_l >> sh) ^ pe) & 1u);
      const unsigned R3 = (F0 >> 1) | ((((unsigned)(we_r >> sh) << 31) ^ pe) & 0x80000000u);
      const unsigned c = __popc((L0 ^ xl) & (F0 ^ xm) & (F1 ^ xr)) + __popc((F0 ^ xl) & (F1 ^ xm) & (F2 ^ xr)) +
                         __popc((F1 ^ xl) & (F2 ^ xm) & (F3 ^ xr)) + __popc((F2 ^ xl) & (F3 ^ xm) & (R3 ^ xr));
      if (lane < 8 && c) atomicAdd(&T.C[bi*8 + k], c);
      const unsigned head = ((unsigned)bi << 26) | tcode;
      if (J0 | J1 | J2 | J3) {
        // queue order: by lane, then by row (any order will do; one prefix per lane, not per mask)
        int q = qn + __popc(J0 & lt) + __popc(J1 & lt) + __popc(J2 & lt) + __popc(J3 & lt);
        const unsigned e = head | (unsigned)(4*lane + 1);
        if (n0) s_queue[q++] = e;
        if (n1) s_queue[q++] = e + 1;
        if (n2) s_queue[q++] = e + 2;
        if (n3) s_queue[q] = e + 3;
        qn += __popc(J0) + __popc(J1) + __popc(J2) + __popc(J3);
      }
      if (edge_l | edge_r) {                     // warp-uniform, two tiles of the whole launch
        const unsigned je = __ballot_sync(FULL, ne != 0) & ((edge_l ? 1u : 0u) | (edge_r ? 0x80000000u : 0u));
        if ((je >> lane) & 1) s_queue[qn + __popc(je & lt)] = head | (lane == 0 ? 0u : (unsigned)(ST_ROWS + 1));
        qn += __popc(je);
      }
      const bool last = bi + 1 == n_br && tile + n_warps >= n_tiles;   // this warp's last step
      while (qn >= 32 || (last && qn > 0)) {
        const int n_take = qn < 32 ? qn : 32;
        st_drain(pv, state, W, T, s_queue + (qn - n_take), n_take, first_s0, ord_stride, n_adds);
        qn -= n_take;
      }
    }
  }
  __syncthreads();
  // integer-only global accumulation; columns are summed here
  for (int i = threadIdx.x; i < n_br*24; i += ST_THREADS) {
    unsigned long long v = 0;
#if ST_SPLIT_
    v = (unsigned long long)T.S[i];
    for (int cix = 0; cix < (1 << col_shift); ++cix) {
      const unsigned *pp = T.P + 3*(((size_t)i << col_shift) + cix);
      v += (unsigned long long)pp[0] + ((unsigned long long)pp[1] << 14) + ((unsigned long long)pp[2] << 28);
    }
#else
    for (int cix = 0; cix < (1 << col_shift); ++cix) v += (unsigned long long)T.S[((size_t)i << col_shift) + cix];
#endif
    if (v) {
      const int bi = i/24, cls = i%24;
      atomicAdd(acc + bi*SS_ACC + 32 + cls, v & 0xffffffffull);
      atomicAdd(acc + bi*SS_ACC + 56 + cls, (unsigned long long)((long long)v >> 32));
    }
  }
  for (int i = threadIdx.x; i < n_br*8; i += ST_THREADS) {
    const int bi = i >> 3, t = i & 7;
    if (T.J[i]) atomicAdd(acc + bi*SS_ACC + 8 + t, (unsigned long long)T.J[i]);
    if (T.C[i]) atomicAdd(acc + bi*SS_ACC + 16 + t, (unsigned long long)T.C[i]);
  }
}

// D[b][t] = (row-merged partial sums + event-time sums with their signs) + count*tot_time[b];
// J[b][t] = count.  Row 0 (root) is zero.  Event class (trip_before, which): which = 0 / 1 / 2 =
// the right / middle / left row of the triple jumped, i.e. bit 1 / 2 / 4 of the triplet flips;
// the time is added to D[trip_before] and taken from D[trip_before ^ flip].
__global__ void
suffstats_fold_kernel(const unsigned long long *acc, const double *branch, int n_nodes,
                      double *J, double *D) {
  const int i = blockIdx.x*blockDim.x + threadIdx.x;
  if (i >= n_nodes*8) return;
  const int b = i >> 3, t = i & 7;
  if (b == 0) { J[i] = 0.0; D[i] = 0.0; return; }
  const unsigned long long *a = acc + (b - 1)*SS_ACC;
  long long lo = (long long)a[t], hi = (long long)a[24 + t];
#pragma unroll
  for (int w = 0; w < 3; ++w) {
    const int out = t*3 + w, in = (t ^ (1 << w))*3 + w;
    lo += (long long)a[32 + out] - (long long)a[32 + in];
    hi += (long long)a[56 + out] - (long long)a[56 + in];
  }
  D[i] = ((double)hi*4294967296.0 + (double)lo)*(branch[b]/SS_FIX) + (double)a[16 + t]*branch[b];
  J[i] = (double)a[8 + t];
}

struct Rates8 { double r[8]; };

// log_likelihood (ParamEstimation.cpp:131-143) and get_gradient (:147-194) on J/D
// collapsed over branches (:299-315).  out = [llh, g0..g7].
__global__ void
llh_grad_kernel(const double *Jb, const double *Db, int n_nodes, Rates8 R, double *out) {
  if (threadIdx.x != 0) return;
  double J[8], D[8];
  for (int t = 0; t < 8; ++t) { J[t] = 0.0; D[t] = 0.0; }
  for (int b = 1; b < n_nodes; ++b)
    for (int t = 0; t < 8; ++t) { J[t] += Jb[b*8 + t]; D[t] += Db[b*8 + t]; }
  double ll = 0.0;
  for (int t = 0; t < 8; ++t) ll += J[t]*log(R.r[t]) - D[t]*R.r[t];
  const double *r = R.r;
  const double f = J[7] - D[7]*r[7];
  double g[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  g[0] = J[0] - D[0]*r[0] + f;
  g[2] = J[2] - D[2]*r[2] - f;
  g[1] = J[1] + J[4] - (D[1] + D[4])*r[1] - 2*f;
  g[4] = g[1];
  g[3] = J[3] + J[6] - (D[3] + D[6])*r[3] + 2*f;
  g[6] = g[3];
  g[5] = J[5] - D[5]*r[5] + f;
  out[0] = ll;
  for (int t = 0; t < 8; ++t) out[1 + t] = g[t];
}

__global__ void
axpy_kernel(double *Js, double *Ds, const double *J, const double *D, int n) {
  const int i = blockIdx.x*blockDim.x + threadIdx.x;
  if (i < n) { Js[i] += J[i]; Ds[i] += D[i]; }
}

// scale_jump_times (ParamEstimation.cpp:369-380): jump *= new_branch/tot_time.
__global__ void __launch_bounds__(256)
scale_kernel(double *jmp, const uint64_t *jbase, const uint32_t *off, uint64_t pitch,
             uint64_t n_sites, const double *scale) {
  const int bi = blockIdx.y;
  const uint64_t n = off[(uint64_t)bi*pitch + n_sites];
  double *p = jmp + jbase[bi];
  const double sc = scale[bi + 1];
  for (uint64_t k = (uint64_t)blockIdx.x*blockDim.x + threadIdx.x; k < n;
       k += (uint64_t)gridDim.x*blockDim.x)
    p[k] *= sc;
}

// Host exchange format (node-major, all local sites) -> device layout: CSR offsets of
// the owned sites [own_lo, own_lo + n_own], state bit sets of every local site.
__global__ void __launch_bounds__(256)
import_kernel(const uint8_t *h_init, const unsigned long long *h_off, uint64_t n_sites,
              int n_nodes, uint64_t own_lo, uint64_t n_own, uint32_t *off, uint64_t pitch,
              uint8_t *state, int W) {
  const uint64_t s = (uint64_t)blockIdx.x*blockDim.x + threadIdx.x;
  if (s <= n_own)
    for (int b = 1; b < n_nodes; ++b) {
      const unsigned long long *o = h_off + (uint64_t)b*n_sites + own_lo;
      off[(uint64_t)(b - 1)*pitch + s] = (uint32_t)(o[s] - o[0]);
    }
  if (s >= n_sites) return;
  for (int w = 0; w < W; ++w) {
    unsigned bits = 0;
    for (int k = 0; k < 8; ++k) {
      const int v = w*8 + k;
      if (v >= 1 && v < n_nodes) bits |= (unsigned)(h_init[(uint64_t)v*n_sites + s] != 0) << k;
    }
    state[s*W + w] = (uint8_t)bits;
  }
}

// Device layout -> host exchange format.  jprefix[bi] = jumps held by branches < bi
// (halo rows included); within a branch the order is halo-low rows, owned rows,
// halo-high rows, i.e. site order.
__global__ void __launch_bounds__(256)
export_kernel(const PathView pv, const uint8_t *state, int W, uint64_t n_sites, int n_nodes,
              const uint64_t *jprefix, uint8_t *h_init, unsigned long long *h_off) {
  const uint64_t s = (uint64_t)blockIdx.x*blockDim.x + threadIdx.x;
  if (s >= n_sites) return;
  h_init[s] = 0; h_off[s] = 0;
  const uint64_t n_own = pv.own_hi - pv.own_lo;
  for (int b = 1; b < n_nodes; ++b) {
    const int bi = b - 1;
    h_init[(uint64_t)b*n_sites + s] = (uint8_t)state_bit(state, W, s, b);
    uint64_t lo_tot = 0, before = 0;
    for (uint64_t k = 0; k < pv.own_lo; ++k) {
      if (k < s) before += pv.hcnt[k*pv.n_br + bi];
      lo_tot += pv.hcnt[k*pv.n_br + bi];
    }
    uint64_t o;
    if (s < pv.own_lo) o = before;
    else if (s < pv.own_hi) o = lo_tot + pv.off[(uint64_t)bi*pv.off_pitch + (s - pv.own_lo)];
    else {
      o = lo_tot + pv.off[(uint64_t)bi*pv.off_pitch + n_own];
      for (uint64_t k = pv.own_hi; k < s; ++k) o += pv.hcnt[(HALO_W + (k - pv.own_hi))*pv.n_br + bi];
    }
    h_off[(uint64_t)b*n_sites + s] = jprefix[bi] + o;
  }
  if (s == 0) h_off[(uint64_t)n_nodes*n_sites] = jprefix[n_nodes - 1];
}

__global__ void
scale_halo_kernel(double *hjmp, int n_br, int halo_k, const double *scale) {
  const int row = blockIdx.x;                     // slot*n_br + bi
  const double sc = scale[row % n_br + 1];
  for (int k = threadIdx.x; k < halo_k; k += blockDim.x) hjmp[(size_t)row*halo_k + k] *= sc;
}

}  // namespace

size_t
suffstats_scratch_bytes(int n_nodes, int) { return (size_t)(n_nodes - 1)*SS_ACC*8; }

namespace {

void
launch_suffstats_generic(const PathView &pv, const uint8_t *state, int W, int n_br, const double *branch,
                         uint64_t lo, uint64_t hi, unsigned long long *scratch, cudaStream_t st,
                         int sm_count) {
  const uint64_t n_chunks = (hi - lo + 32*SS_ITERS - 1)/(32*SS_ITERS);
  const uint64_t want = (n_chunks + SS_THREADS/32 - 1)/(SS_THREADS/32);
  const unsigned blocks = (unsigned)(want < (uint64_t)sm_count*8 ? want : (uint64_t)sm_count*8);
  const size_t smem = (size_t)n_br*8*(8 + 4 + 4) + (SS_THREADS/32)*SS_QCAP*sizeof(unsigned);
  cudaFuncSetAttribute(suffstats_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  suffstats_kernel<<<blocks, SS_THREADS, smem, st>>>(pv, state, W, n_br, branch, lo, hi, scratch);
}

}  // namespace

__global__ void
acc_add_kernel(unsigned long long *dst, const unsigned long long *src, int n) {
  const int i = blockIdx.x*blockDim.x + threadIdx.x;
  if (i < n) dst[i] += src[i];
}

void
launch_suffstats_fold(const unsigned long long *scratch, const TreeDev &tree, double *J, double *D, cudaStream_t st) {
  suffstats_fold_kernel<<<(tree.n_nodes*8 + 127)/128, 128, 0, st>>>(scratch, tree.branch, tree.n_nodes, J, D);
}

void
launch_suffstats_add(unsigned long long *dst, const unsigned long long *src, int n_nodes, cudaStream_t st) {
  const int n = (n_nodes - 1)*SS_ACC;
  acc_add_kernel<<<(n + 127)/128, 128, 0, st>>>(dst, src, n);
}

// The integer accumulators of the pass are ADDED to scratch (cleared first when clear is set); fold turns
// them into J / D.  All-integer => sums over passes and over shards are exact and order-independent.
// Returns the number of kernels launched.
int
launch_suffstats(const PathView &pv, const uint8_t *state, int W, const TreeDev &tree,
                 uint64_t, uint64_t lo, uint64_t hi, double *J, double *D,
                 unsigned long long *scratch, cudaStream_t st, int sm_count, bool clear, bool fold) {
  const int n_br = tree.n_nodes - 1;
  int launches = fold ? 1 : 0;
  if (clear) cudaMemsetAsync(scratch, 0, (size_t)n_br*SS_ACC*8, st);
  if (hi > lo) {
    // rows [r0, r0 + 128 n_tiles): interior sites whose neighbours are owned, offsets r0-1 ..
    // r0 + 128 n_tiles + 1 inside the CSR; the remainder on either side goes to the generic kernel
    const uint64_t n_own = pv.own_hi - pv.own_lo;
    const uint64_t first = lo > pv.own_lo ? lo - pv.own_lo : 0;
    uint64_t r0 = (first + 3) & ~3ull;
    if (r0 < 4) r0 = 4;
    const uint64_t last = hi - pv.own_lo < n_own - 1 ? hi - pv.own_lo : n_own - 1;   // rows < last
#ifdef SS_NO_TILE
    const uint64_t n_tiles = 0;
#else
    const uint64_t n_tiles = (W <= 8 && n_own >= 2 && last > r0) ? (last - r0)/ST_ROWS : 0;
#endif
    if (n_tiles == 0) { launch_suffstats_generic(pv, state, W, n_br, tree.branch, lo, hi, scratch, st, sm_count); ++launches; }
    else {
      const uint64_t t_lo = pv.own_lo + r0, t_hi = t_lo + n_tiles*ST_ROWS;
      const uint64_t want = (n_tiles + ST_THREADS/32 - 1)/(ST_THREADS/32);
      const uint64_t wave = (uint64_t)sm_count*ST_MINB_*ST_WAVES_;
      const unsigned blocks = (unsigned)(want < wave ? want : wave);
      // bin columns: as many as fit in ~24 KB (fewer CAS retries), at least 2
      int col_shift = 5;
#if ST_SPLIT_
      while (col_shift > 0 && ((size_t)n_br*24*12 << col_shift) > ST_BINKB_*1024) --col_shift;
      const size_t smem = ((size_t)n_br*24*12 << col_shift) + (size_t)n_br*24*8 + (size_t)n_br*8 + (size_t)n_br*8*(4 + 4) +
#else
      while (col_shift > 1 && ((size_t)n_br*24*8 << col_shift) > ST_BINKB_*1024) --col_shift;
      const size_t smem = ((size_t)n_br*24*8 << col_shift) + (size_t)n_br*8 + (size_t)n_br*8*(4 + 4) +
#endif
                          (ST_THREADS/32)*ST_QCAP*sizeof(unsigned);
      auto kernel = W <= 4 ? suffstats_tile_kernel<uint32_t> : suffstats_tile_kernel<uint64_t>;
      cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      kernel<<<blocks, ST_THREADS, smem, st>>>(pv, state, W, n_br, tree.branch, r0, n_tiles, col_shift, scratch);
      ++launches;
      if (t_lo > lo) { launch_suffstats_generic(pv, state, W, n_br, tree.branch, lo, t_lo, scratch, st, sm_count); ++launches; }
      if (hi > t_hi) { launch_suffstats_generic(pv, state, W, n_br, tree.branch, t_hi, hi, scratch, st, sm_count); ++launches; }
    }
  }
  if (fold) launch_suffstats_fold(scratch, tree, J, D, st);
  return launches;
}

void
launch_llh_grad(const double *J, const double *D, int n_nodes, const double *rates,
                double *out9, cudaStream_t st) {
  Rates8 R;
  for (int i = 0; i < 8; ++i) R.r[i] = rates[i];
  llh_grad_kernel<<<1, 32, 0, st>>>(J, D, n_nodes, R, out9);
}

void
launch_axpy_stats(double *Js, double *Ds, const double *J, const double *D, int n,
                  cudaStream_t st) {
  axpy_kernel<<<(n + 127)/128, 128, 0, st>>>(Js, Ds, J, D, n);
}

void
launch_scale_jumps(double *jmp, const uint64_t *jbase, const uint32_t *off, uint64_t pitch,
                   uint64_t n_own, int n_branches, const double *scale, double *hjmp, int halo_k,
                   cudaStream_t st) {
  dim3 grid(592, n_branches);
  scale_kernel<<<grid, 256, 0, st>>>(jmp, jbase, off, pitch, n_own, scale);
  scale_halo_kernel<<<2*HALO_W*n_branches, 64, 0, st>>>(hjmp, n_branches, halo_k, scale);
}

void
launch_import(const uint8_t *h_init, const unsigned long long *h_off, uint64_t n_sites,
              int n_nodes, uint64_t own_lo, uint64_t n_own, uint32_t *off, uint64_t pitch,
              uint8_t *state, int W, cudaStream_t st) {
  import_kernel<<<(unsigned)((n_sites + 256)/256), 256, 0, st>>>(h_init, h_off, n_sites, n_nodes,
                                                                 own_lo, n_own, off, pitch, state, W);
}

void
launch_export(const PathView &pv, const uint8_t *state, int W, uint64_t n_sites, int n_nodes,
              const uint64_t *jprefix, uint8_t *h_init, unsigned long long *h_off,
              cudaStream_t st) {
  export_kernel<<<(unsigned)((n_sites + 255)/256), 256, 0, st>>>(pv, state, W, n_sites, n_nodes,
                                                                 jprefix, h_init, h_off);
}

}  // namespace epv
