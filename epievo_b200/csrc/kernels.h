// kernels.h -- host-visible launch interface of the CUDA kernels (internal to the .so).
#pragma once
#include "common.cuh"

namespace epv {

constexpr int MH_THREADS = 128;                 // update kernel: groups (of 3 sites) per CTA of the small launches (edge
                                                // groups, overflow list) and of trees wider than 64 nodes; the main
                                                // launch runs one large CTA per SM (mh_threads(), kernels_mh.cu)
constexpr int RB_THREADS = 256;                 // rebuild kernel: 4 CSR rows per thread
constexpr int RB_TILE_ROWS = 4*RB_THREADS;

struct PassArgs {
  ModelParams mp;
  PathView in;                // snapshot read by every update of this colour pass
  uint32_t *off_out;          // rebuilt CSR (ping-pong partner), same pitch / jbase
  double *jmp_out;
  const uint64_t *jcap;       // [n_branches] capacity of each branch's jump slice
  unsigned long long *jneed;  // [n_branches] on overflow: the total the rebuilt branch needs (atomicMax)
  uint8_t *state;             // in place: only the updated site's own bits change
  int W;
  TreeDev tree;
  double *tri_llh;            // in place: an update writes [s-1, s+1]
  uint64_t n_sites;           // local sites (shard + halos)
  uint64_t upd_lo, upd_hi;    // local site range this context updates
  int64_t global0;            // global index of local site 0
  uint64_t n_global;
  int colour;
  int q_slots;                // internal non-root nodes of the tree (q kept in shared memory)
  int coop;                   // warp-cooperative jump-time sampling (pays when jumps are sparse)
  uint64_t seed;
  uint32_t sweep;
  // single-pass chained scan (decoupled look-back) state
  unsigned long long *desc;   // [n_branches * n_tiles]
  uint32_t n_tiles;
  uint32_t epoch;
  uint32_t *ticket;
  uint32_t ticket_base;
  // overflow scratch for proposals longer than the per-thread buffer
  double *scratch;
  uint32_t *scratch_ctr;
  uint32_t scratch_slots, scratch_cap;
  // staging of the accepted proposals' rows between the update and the rebuild kernel
  uint64_t n_groups;          // groups of 3 sites covering the local sites
  // groups one launch works on: index i < g_count maps to group g_first + i (+ g_skip when i >= g_split)
  uint64_t g_first, g_split, g_skip, g_count;
  int reverse;                // test-only (EPIEVO_B200_REVERSE): thread i takes group g_count-1-i -- another schedule, same chain
  uint32_t *ovf_list;         // [1 + n_groups]: count, then the groups the main update launch deferred
  int list_mode;              // the launch works on ovf_list instead of a group range
  uint32_t *stage_off;        // [n_groups] offset into stage, or rejected / empty marker
  double *stage;              // blobs: ceil(n_br/4) words of u16 jump counts + jump times
  unsigned long long *stage_ctr;
  unsigned long long stage_cap;
  unsigned *err;
  unsigned long long *n_accept;
};

int launch_mh_pass(const PassArgs &a, cudaStream_t st);       // update kernel: interior groups, then edge groups
int launch_mh_interior(const PassArgs &a, cudaStream_t st);   // groups whose sites s-2..s+2 are all owned (+ their overflow launch)
int launch_mh_edge(const PassArgs &a, cudaStream_t st);       // groups next to a shard's ends (halo-aware)
void launch_csr_rebuild(const PassArgs &a, cudaStream_t st);   // re-compaction kernel
size_t mh_pass_smem_bytes(int n_nodes);

void launch_tri_llh(const ModelParams &mp, const PathView &pv, const uint8_t *state, int W,
                    const TreeDev &tree, uint64_t n_sites, double *tri_llh,
                    cudaStream_t st);

struct DebugOut {
  unsigned long long *n_intervals, *seg_off, *trip0, *trip1;
  double *rate0, *rate1, *len, *p, *q, *root_p0, *logp;
  unsigned long long cap;
};
void launch_site_debug(const ModelParams &mp, const PathView &pv, const uint8_t *state,
                       int W, const TreeDev &tree, uint64_t site, const DebugOut &out,
                       cudaStream_t st);
void launch_end_cond(double r0, double r1, int a, int b, double T, uint64_t seed,
                     uint64_t n, uint64_t cap, uint32_t *n_jumps, double *jumps,
                     cudaStream_t st);

// J/D sufficient statistics: acc = [n_branches][24] doubles per context
// (8 D sums, 8 J counts, 8 event-free row counts are folded on device).
int launch_suffstats(const PathView &pv, const uint8_t *state, int W, const TreeDev &tree,
                     uint64_t n_sites, uint64_t lo, uint64_t hi, double *J, double *D,
                     unsigned long long *scratch, cudaStream_t st, int sm_count, bool clear = true,
                     bool fold = true);   // returns kernels launched
void launch_suffstats_fold(const unsigned long long *scratch, const TreeDev &tree, double *J, double *D, cudaStream_t st);
void launch_suffstats_add(unsigned long long *dst, const unsigned long long *src, int n_nodes, cudaStream_t st);
size_t suffstats_scratch_bytes(int n_nodes, int sm_count);
void launch_llh_grad(const double *J, const double *D, int n_nodes, const double *rates_dev,
                     double *out9, cudaStream_t st);
void launch_axpy_stats(double *J_sum, double *D_sum, const double *J, const double *D, int n,
                       cudaStream_t st);

void launch_scale_jumps(double *jmp, const uint64_t *jbase, const uint32_t *off,
                        uint64_t off_pitch, uint64_t n_own, int n_branches,
                        const double *scale, double *hjmp, int halo_k, cudaStream_t st);

// host exchange format <-> device layout
void launch_import(const uint8_t *h_init, const unsigned long long *h_off,
                   uint64_t n_sites, int n_nodes, uint64_t own_lo, uint64_t n_own,
                   uint32_t *off, uint64_t off_pitch, uint8_t *state, int W, cudaStream_t st);
void launch_export(const PathView &pv, const uint8_t *state, int W, uint64_t n_sites,
                   int n_nodes, const uint64_t *jprefix, uint8_t *h_init,
                   unsigned long long *h_off, cudaStream_t st);

// compact host exchange format (kernels_compact.cu)
void launch_compact_totals(const uint16_t *code, uint64_t n_sites, int n_br, uint64_t own_lo,
                           uint64_t own_hi, unsigned long long *totals, cudaStream_t st);
void launch_compact_import(const uint16_t *code, uint64_t n_sites, int n_nodes, uint64_t own_lo,
                           uint64_t n_own, uint32_t *off, uint64_t pitch, uint8_t *state, int W,
                           unsigned long long *desc, uint32_t n_tiles, uint32_t epoch,
                           uint32_t *ticket, uint32_t ticket_base, cudaStream_t st);
void launch_compact_export(const PathView &pv, const uint8_t *state, int W, uint64_t n_sites,
                           int n_nodes, uint16_t *code, unsigned *err, cudaStream_t st);

// ---- site-independent E-step of epievo_initialization (kernels_indep.cu) -----------------
struct IndepTab { double P[4], J0[4], J1[4], D0[4], D1[4]; };   // per node: 2x2 matrices, row-major (start, end)
void launch_indep_expect(const PathView &pv, const uint8_t *state, int W, const TreeDev &tree,
                         const IndepTab *tab, double pi0, double *acc, int sm_count, cudaStream_t st);
void launch_indep_stats(const PathView &pv, const uint8_t *state, int W, const TreeDev &tree, double *acc,
                        int sm_count, cudaStream_t st);
void launch_indep_sample(const PassArgs &a, const IndepTab *tab, double r0, double r1, cudaStream_t st);

// ---- halo exchange (multi-GPU): one boundary site per side per colour pass -------------
// Record (8-byte words): [0..2] tri_llh[s-1], [s], [s+1]; [3] global site index;
// [4 .. 4+SW) state bytes; then n_br u16 jump counts packed in CW words; then
// n_br*halo_k jump times (halo_k = pv.halo_k, the same on every shard of a run).
constexpr uint32_t STG_REJECTED = 0xffffffffu, STG_EMPTY = 0xfffffffeu;   // stage_off markers
struct HaloArgs {
  PathView pv;
  uint8_t *state; int W;
  double *tri_llh;
  uint64_t n_sites; int64_t global0;
  uint16_t *hcnt; double *hjmp;        // writable halo store (unpack)
  unsigned *err;
  unsigned *hmax;                      // largest jump count of a packed boundary row seen so far (atomicMax)
  // staged variant of the pack: pv is the CSR the colour pass READ; the boundary site's new rows are
  // taken from the proposal staging blob when its update was accepted (before the CSR is rebuilt)
  const uint32_t *stage_off;
  const double *stage;
};
size_t halo_record_words(int n_nodes, int halo_k);
// site_lo / site_hi: local index of the owned boundary site to pack (or -1: side absent)
void launch_halo_pack(const HaloArgs &a, int64_t site_lo, unsigned long long *rec_lo,
                      int64_t site_hi, unsigned long long *rec_hi, cudaStream_t st, bool staged = false);
// site_lo / site_hi: local index of the halo site each record must describe (or -1)
void launch_halo_unpack(const HaloArgs &a, int64_t site_lo, const unsigned long long *rec_lo,
                        int64_t site_hi, const unsigned long long *rec_hi, cudaStream_t st);

}  // namespace epv
