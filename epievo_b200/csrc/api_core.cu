// api_core.cu -- context lifetime, tree/model/shard setup, path upload/download.
#include <cmath>
#include <cstdio>
#include <cstring>

#include "../../include/epievo_b200.h"
#include "context.h"

namespace epv {

int fail(epv_ctx *c, const std::string &msg) { if (c) c->error = msg; return 1; }

int check_cuda(epv_ctx *c, cudaError_t e, const char *what) {
  if (e == cudaSuccess) return 0;
  return fail(c, std::string("CUDA error: ") + cudaGetErrorString(e) + " in " + what);
}

void free_paths(epv_ctx *c) {
  for (int k = 0; k < 2; ++k) {
    cudaFree(c->d_off[k]); c->d_off[k] = nullptr;
    cudaFree(c->d_jmp[k]); c->d_jmp[k] = nullptr;
  }
  cudaFree(c->d_jbase); cudaFree(c->d_jcap); cudaFree(c->d_state); cudaFree(c->d_tri);
  cudaFree(c->d_jneed); c->d_jneed = nullptr;
  cudaFree(c->d_desc); cudaFree(c->d_hcnt); cudaFree(c->d_hjmp);
  c->halo_k = HALO_K_MIN;
  cudaFree(c->d_stage_init); cudaFree(c->d_stage_off); cudaFree(c->d_stg_off); cudaFree(c->d_stg);
  cudaFree(c->d_ovf_list); c->d_ovf_list = nullptr;
  c->d_stg_off = nullptr; c->d_stg = nullptr;
  c->d_hcnt = nullptr; c->d_hjmp = nullptr; c->d_stage_init = nullptr; c->d_stage_off = nullptr;
  c->d_jbase = c->d_jcap = nullptr; c->d_state = nullptr; c->d_tri = nullptr;
  c->d_desc = nullptr;
  c->n_sites = 0; c->tri_valid = false; c->stats_valid = false; c->h_jcap.clear();
  c->hk_pending = false;
}

int check_device_flags(epv_ctx *c) {
  if (int r = housekeeping_take(c)) return r;      // (a record of the last sweep that nobody looked at yet)
  unsigned flags = 0;
  EPV_CUDA(c, cudaMemcpyAsync(c->h_mail, c->d_err, sizeof(unsigned), cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  flags = *reinterpret_cast<unsigned *>(c->h_mail);
  if (!flags) return 0;
  std::string m = "device-side failure:";
  if (flags & ERR_JUMP_CAPACITY) m += " jump storage capacity exceeded;";
  if (flags & ERR_SCRATCH_SLOTS) m += " too many long proposals in one pass;";
  if (flags & ERR_SCRATCH_SIZE) m += " a proposed path exceeds the per-site scratch capacity;";
  if (flags & ERR_SAMPLER_STUCK) m += " end-conditioned sampler did not terminate;";
  if (flags & ERR_BAD_INPUT) m += " malformed input;";
  if (flags & ERR_STAGE_CAPACITY) m += " proposal staging buffer exceeded;";
  if (flags & ERR_HALO_CAPACITY) m += " a shard-boundary row has more jumps than a halo record holds;";
  if (flags & ERR_HALO_MISMATCH) m += " a halo record arrived for the wrong site (driver / colour mismatch);";
  return fail(c, m + " paths on the device are no longer valid");
}

}  // namespace epv

using namespace epv;

extern "C" {

int epv_create(epv_ctx **out, int device) {
  if (!out) return 1;
  *out = nullptr;
  int n_dev = 0;
  if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0 || device < 0 || device >= n_dev) {
    std::fprintf(stderr, "epievo_b200: no usable CUDA device %d (there is no CPU fallback)\n", device);
    return 1;
  }
  epv_ctx *c = new epv_ctx;
  c->device = device;
  bool ok = cudaSetDevice(device) == cudaSuccess
    && cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking) == cudaSuccess
    && cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device) == cudaSuccess
    && cudaMalloc(&c->d_err, sizeof(unsigned)) == cudaSuccess
    && cudaMalloc(&c->d_hmax, sizeof(unsigned)) == cudaSuccess
    && cudaStreamCreateWithFlags(&c->own_side, cudaStreamNonBlocking) == cudaSuccess
    && cudaEventCreateWithFlags(&c->ev_begin, cudaEventDisableTiming) == cudaSuccess
    && cudaEventCreateWithFlags(&c->ev_edge, cudaEventDisableTiming) == cudaSuccess
    && cudaEventCreateWithFlags(&c->ev_halo, cudaEventDisableTiming) == cudaSuccess
    && cudaMalloc(&c->d_nacc, sizeof(unsigned long long)) == cudaSuccess
    && cudaMalloc(&c->d_ticket, sizeof(uint32_t)) == cudaSuccess
    && cudaMalloc(&c->d_scratch_ctr, sizeof(uint32_t)) == cudaSuccess
    && cudaMalloc(&c->d_stg_ctr, sizeof(unsigned long long)) == cudaSuccess
    && cudaMalloc(&c->d_out9, 9*sizeof(double)) == cudaSuccess
    && cudaMalloc(&c->d_mail, 4096*sizeof(unsigned long long)) == cudaSuccess
    && cudaMallocHost(&c->h_mail, 4096*sizeof(unsigned long long)) == cudaSuccess
    && cudaMallocHost(&c->h_hk, 4096*sizeof(uint32_t)) == cudaSuccess
    && cudaEventCreateWithFlags(&c->ev_hk, cudaEventDisableTiming) == cudaSuccess
    && cudaMalloc(&c->d_scratch, (size_t)c->scratch_slots*c->scratch_cap*sizeof(double)) == cudaSuccess
    && cudaEventCreate(&c->ev0) == cudaSuccess && cudaEventCreate(&c->ev1) == cudaSuccess;
  for (int k = 0; ok && k < 3*epv_ctx::EV_RING; ++k)
    ok = cudaEventCreate(&c->ev_ring[k]) == cudaSuccess;
  if (ok) {
    c->stream = c->own_stream;
    c->side_stream = c->own_side;
    cudaMemset(c->d_err, 0, sizeof(unsigned));
    cudaMemset(c->d_hmax, 0, sizeof(unsigned));
    cudaMemset(c->d_nacc, 0, sizeof(unsigned long long));
    cudaMemset(c->d_ticket, 0, sizeof(uint32_t));
    cudaMemset(c->d_scratch_ctr, 0, sizeof(uint32_t));
  }
  if (!ok) {
    std::fprintf(stderr, "epievo_b200: context setup failed: %s\n", cudaGetErrorString(cudaGetLastError()));
    delete c;
    return 1;
  }
  *out = c;
  return 0;
}

void epv_destroy(epv_ctx *c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  free_paths(c);
  sim_release(c);
  cudaFree(c->d_parent); cudaFree(c->d_subtree); cudaFree(c->d_branch);
  cudaStreamSynchronize(c->side_stream);
  cudaEventDestroy(c->ev_begin); cudaEventDestroy(c->ev_edge); cudaEventDestroy(c->ev_halo);
  cudaStreamDestroy(c->own_side);
  cudaFree(c->d_hmax);
  cudaFree(c->d_err); cudaFree(c->d_nacc); cudaFree(c->d_ticket); cudaFree(c->d_scratch_ctr);
  cudaFree(c->d_stg_ctr); cudaFree(c->d_itab); cudaFree(c->d_iacc);
  cudaFree(c->d_scratch); cudaFree(c->d_out9); cudaFree(c->d_mail); cudaFreeHost(c->h_mail);
  cudaFreeHost(c->h_hk); if (c->ev_hk) cudaEventDestroy(c->ev_hk);
  cudaFree(c->d_J); cudaFree(c->d_D); cudaFree(c->d_Jsum); cudaFree(c->d_Dsum); cudaFree(c->d_ss);
  cudaEventDestroy(c->ev0); cudaEventDestroy(c->ev1);
  for (int k = 0; k < 3*epv_ctx::EV_RING; ++k) if (c->ev_ring[k]) cudaEventDestroy(c->ev_ring[k]);
  cudaStreamDestroy(c->own_stream);
  delete c;
}

const char *epv_last_error(const epv_ctx *c) { return c ? c->error.c_str() : "null context"; }

int epv_set_stream(epv_ctx *c, void *s) {
  if (!c) return 1;
  cudaStreamSynchronize(c->stream);
  c->stream = s ? (cudaStream_t)s : c->own_stream;
  return 0;
}

int epv_synchronize(epv_ctx *c) {
  if (!c) return 1;
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->side_stream));
  return c->n_sites ? housekeeping_take(c) : 0;
}

int epv_set_side_stream(epv_ctx *c, void *s) {
  if (!c) return 1;
  cudaStreamSynchronize(c->side_stream);
  c->side_stream = s ? (cudaStream_t)s : c->own_side;
  return 0;
}

int epv_set_tree(epv_ctx *c, uint64_t n_nodes, const uint64_t *subtree_sizes,
                 const uint64_t *parent_ids, const double *branches) {
  if (!c) return 1;
  // the CSR rebuild keeps one u32 per (branch, thread) in shared memory: (n_nodes-1) KB <= 227 KB
  if (n_nodes < 2 || n_nodes > 224) return fail(c, "tree must have 2..224 nodes");
  for (uint64_t v = 1; v < n_nodes; ++v) {
    if (parent_ids[v] >= v) return fail(c, "tree is not in preorder (parent_ids[v] >= v)");
    if (!(branches[v] > 0.0)) return fail(c, "branch lengths must be positive");
  }
  if (subtree_sizes[0] != n_nodes) return fail(c, "subtree_sizes[0] must equal n_nodes");
  cudaSetDevice(c->device);
  if ((int)n_nodes != c->n_nodes) {
    if (c->n_sites) return fail(c, "cannot change the node count while paths are loaded");
    cudaFree(c->d_parent); cudaFree(c->d_subtree); cudaFree(c->d_branch);
    cudaFree(c->d_J); cudaFree(c->d_D); cudaFree(c->d_Jsum); cudaFree(c->d_Dsum); cudaFree(c->d_ss);
    EPV_CUDA(c, cudaMalloc(&c->d_parent, n_nodes*sizeof(int)));
    EPV_CUDA(c, cudaMalloc(&c->d_subtree, n_nodes*sizeof(int)));
    EPV_CUDA(c, cudaMalloc(&c->d_branch, n_nodes*sizeof(double)));
    EPV_CUDA(c, cudaMalloc(&c->d_J, n_nodes*8*sizeof(double)));
    EPV_CUDA(c, cudaMalloc(&c->d_D, n_nodes*8*sizeof(double)));
    EPV_CUDA(c, cudaMalloc(&c->d_Jsum, n_nodes*8*sizeof(double)));
    EPV_CUDA(c, cudaMalloc(&c->d_Dsum, n_nodes*8*sizeof(double)));
    EPV_CUDA(c, cudaMalloc(&c->d_ss, suffstats_scratch_bytes((int)n_nodes, c->sm_count)));
  }
  c->n_nodes = (int)n_nodes;
  c->h_parent.assign(n_nodes, 0); c->h_subtree.assign(n_nodes, 0);
  c->h_branch.assign(branches, branches + n_nodes);
  c->h_branch[0] = 0.0;
  for (uint64_t v = 0; v < n_nodes; ++v) {
    c->h_parent[v] = (int)parent_ids[v]; c->h_subtree[v] = (int)subtree_sizes[v];
  }
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  EPV_CUDA(c, cudaMemcpy(c->d_parent, c->h_parent.data(), n_nodes*sizeof(int), cudaMemcpyHostToDevice));
  EPV_CUDA(c, cudaMemcpy(c->d_subtree, c->h_subtree.data(), n_nodes*sizeof(int), cudaMemcpyHostToDevice));
  EPV_CUDA(c, cudaMemcpy(c->d_branch, c->h_branch.data(), n_nodes*sizeof(double), cudaMemcpyHostToDevice));
  c->tri_valid = false;
  return 0;
}

int epv_set_model(epv_ctx *c, const double *rates, const double *T) {
  if (!c) return 1;
  for (int i = 0; i < 8; ++i) {
    if (!(rates[i] > 0.0) || !std::isfinite(rates[i])) return fail(c, "triplet rates must be positive and finite");
    c->mp.rate[i] = rates[i];
    c->mp.log_rate[i] = std::log(rates[i]);
  }
  for (int i = 0; i < 4; ++i) c->mp.T[i] = T[i];
  c->have_model = true;
  c->tri_valid = false;
  return 0;
}

int epv_set_shard(epv_ctx *c, uint64_t n_global, uint64_t site_begin, uint64_t site_end,
                  uint64_t halo_lo, uint64_t halo_hi) {
  if (!c) return 1;
  if (c->n_sites) return fail(c, "set the shard before uploading paths");
  if (site_begin % 3) return fail(c, "site_begin must be a multiple of 3 (colour alignment)");
  if (site_end <= site_begin || site_end > n_global || halo_lo > site_begin
      || site_end + halo_hi > n_global)
    return fail(c, "inconsistent shard bounds");
  if ((halo_lo != 0 && halo_lo != HALO_W) || (halo_hi != 0 && halo_hi != HALO_W))
    return fail(c, "halo_lo / halo_hi must be 0 (global end) or 3 sites");
  if ((site_begin > 0) != (halo_lo != 0) || (site_end < n_global) != (halo_hi != 0))
    return fail(c, "a shard needs a 3-site halo on every side that has a neighbour");
  if (site_end - site_begin < 2*HALO_W) return fail(c, "a shard must own at least 6 sites");
  c->have_shard = true;
  c->n_global = n_global; c->site_begin = site_begin; c->site_end = site_end;
  c->halo_lo = halo_lo; c->halo_hi = halo_hi;
  return 0;
}

}  // extern "C"
