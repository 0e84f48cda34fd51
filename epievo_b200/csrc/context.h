// context.h -- the opaque epv_ctx behind the C-ABI (internal).
#pragma once
#include <string>
#include <vector>

#include "kernels.h"

struct epv_ctx {
  int device = 0, sm_count = 148;
  cudaStream_t own_stream = nullptr, stream = nullptr;
  std::string error;

  // tree
  int n_nodes = 0;
  std::vector<int> h_parent, h_subtree;
  std::vector<double> h_branch;
  int *d_parent = nullptr, *d_subtree = nullptr;
  double *d_branch = nullptr;
  // model
  bool have_model = false;
  epv::ModelParams mp;
  // shard
  bool have_shard = false;
  uint64_t n_global = 0, site_begin = 0, site_end = 0, halo_lo = 0, halo_hi = 0;

  // paths
  uint64_t n_sites = 0, off_pitch = 0;
  int W = 0;
  uint32_t *d_off[2] = {nullptr, nullptr};
  double *d_jmp[2] = {nullptr, nullptr};
  uint64_t *d_jbase = nullptr, *d_jcap = nullptr;
  unsigned long long *d_jneed = nullptr;     // [n_br] written by a rebuild that overflowed (kernels.h)
  bool guard = false;                        // thin headroom: check (and recover) after every pass, not every sweep
  uint64_t recoveries = 0;
  std::vector<uint64_t> h_jbase, h_jcap;
  uint64_t jmp_total_cap = 0;
  int cur = 0;
  uint8_t *d_state = nullptr;
  uint8_t *d_stage_init = nullptr;           // staging of the host exchange format
  unsigned long long *d_stage_off = nullptr;
  double *d_tri = nullptr;
  bool tri_valid = false;

  // staging of accepted proposals between the update and the rebuild kernel
  uint32_t *d_stg_off = nullptr;             // [n_groups]
  uint32_t *d_ovf_list = nullptr;            // [1 + n_groups] groups deferred by the main update launch
  double *d_stg = nullptr;
  unsigned long long stg_cap = 0, *d_stg_ctr = nullptr;
  uint64_t n_groups = 0;
  // colour-pass machinery
  unsigned long long *d_desc = nullptr;
  uint32_t n_tiles = 0, epoch = 0, ticket_base = 0;
  double jumps_per_row = 0.0;                // owned rows, refreshed whenever the totals are read
  uint32_t *d_ticket = nullptr, *d_scratch_ctr = nullptr;
  double *d_scratch = nullptr;
  uint32_t scratch_slots = 2048, scratch_cap = 2048;
  unsigned *d_err = nullptr;
  unsigned long long *d_nacc = nullptr;

  // statistics
  epv::IndepTab *d_itab = nullptr;           // site-independent model: per-node tables, [n_nodes*4] sums
  double *d_iacc = nullptr;
  int itab_nodes = 0;
  double *d_J = nullptr, *d_D = nullptr, *d_Jsum = nullptr, *d_Dsum = nullptr, *d_out9 = nullptr;
  unsigned long long *d_ss = nullptr;
  bool stats_valid = false;

  // small pinned host mailbox
  unsigned long long *h_mail = nullptr;      // [64 + n_nodes]
  unsigned long long *d_mail = nullptr;
  // per-sweep housekeeping read a sweep late (branch totals + error word copied asynchronously, api_sampler.cu)
  uint32_t *h_hk = nullptr;                  // pinned, [1 + n_branches]
  cudaEvent_t ev_hk = nullptr;
  bool hk_pending = false;

  // halo store (rows of the <= 3 + 3 halo sites, outside the CSR; see common.cuh)
  uint16_t *d_hcnt = nullptr;               // [2*HALO_W][n_br]
  double *d_hjmp = nullptr;                 // [2*HALO_W][n_br][halo_k]
  int halo_k = epv::HALO_K_MIN;             // jump capacity of one halo row (the same on every shard of a run)
  unsigned *d_hmax = nullptr;               // largest boundary-row jump count seen by the pack kernels
  // second stream of a sharded context: edge groups + halo traffic overlap the interior update
  cudaStream_t side_stream = nullptr, own_side = nullptr;
  cudaEvent_t ev_begin = nullptr, ev_edge = nullptr, ev_halo = nullptr;
  bool split_open = false;                  // between epv_pass_begin and epv_pass_end

  // result of epv_simulate until epv_simulate_fetch (api_sim.cu)
  void *sim = nullptr;

  // instrumentation
  uint64_t launches = 0, mh_launches = 0;
  bool timing = false;
  double mh_ms = 0.0, rebuild_ms = 0.0, last_rebuild_ms = 0.0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  // ring of (start, stop) event pairs around the colour-pass launches; resolved
  // lazily (after the per-sweep stream sync) so timing never stalls the stream
  static constexpr int EV_RING = 96;
  cudaEvent_t ev_ring[3*EV_RING] = {};       // (before update, between, after rebuild)
  int ev_pending = 0;

  epv::PathView view(int which) const {
    epv::PathView v; v.off = d_off[which]; v.jmp = d_jmp[which]; v.jbase = d_jbase;
    v.off_pitch = off_pitch; v.own_lo = halo_lo; v.own_hi = n_sites - halo_hi; v.n_local = n_sites;
    v.hcnt = d_hcnt; v.hjmp = d_hjmp; v.n_br = n_nodes - 1; v.halo_k = halo_k; return v;
  }
  epv::TreeDev tree() const {
    epv::TreeDev t; t.n_nodes = n_nodes; t.parent = d_parent; t.subtree = d_subtree;
    t.branch = d_branch; return t;
  }
  uint64_t n_own() const { return n_sites - halo_lo - halo_hi; }
  uint64_t upd_lo() const { return halo_lo; }
  uint64_t upd_hi() const { return n_sites - halo_hi; }
  int64_t global0() const { return (int64_t)site_begin - (int64_t)halo_lo; }
};

namespace epv {
int fail(epv_ctx *c, const std::string &msg);
int check_cuda(epv_ctx *c, cudaError_t e, const char *what);
void free_paths(epv_ctx *c);
int check_device_flags(epv_ctx *c);
int housekeeping_take(epv_ctx *c);     // api_sampler.cu: look at the last sweep's late housekeeping record, if any
int set_halo_capacity(epv_ctx *c, int k);  // reallocates the halo store (contents kept)
// pieces of the upload path shared with epv_simulate_adopt (api_paths.cu)
int64_t slot_site(const epv_ctx *c, int k);
int check_upload(epv_ctx *c, uint64_t n_sites, uint64_t n_nodes, const double *tot_time);
int ensure_shape(epv_ctx *c, uint64_t n_sites);
int ensure_jumps(epv_ctx *c, const std::vector<uint64_t> &count);
void sim_release(epv_ctx *c);          // frees what epv_simulate left on the device   // reads d_err, turns sticky flags into an error
}
#define EPV_CUDA(c, call) do { if (int _r = epv::check_cuda((c), (call), #call)) return _r; } while (0)
