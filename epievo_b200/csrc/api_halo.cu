// api_halo.cu -- C-ABI of the per-colour halo exchange (multi-GPU site sharding).
#include <algorithm>

#include "../../include/epievo_b200.h"
#include "context.h"

using namespace epv;

namespace {

HaloArgs halo_args(epv_ctx *c) {
  HaloArgs a;
  a.pv = c->view(c->cur);
  a.state = c->d_state; a.W = c->W; a.tri_llh = c->d_tri;
  a.n_sites = c->n_sites; a.global0 = c->global0();
  a.hcnt = c->d_hcnt; a.hjmp = c->d_hjmp; a.err = c->d_err; a.hmax = c->d_hmax;
  a.stage_off = c->d_stg_off; a.stage = c->d_stg;
  return a;
}

int halo_ready(epv_ctx *c, int colour) {
  if (!c) return 1;
  if (!c->n_sites) return fail(c, "no paths loaded");
  if (colour < 0 || colour > 2) return fail(c, "colour must be 0, 1 or 2");
  cudaSetDevice(c->device);
  return 0;
}

}  // namespace

extern "C" {

int epv_halo_record_bytes(epv_ctx *c, uint64_t *bytes) {
  if (!c) return 1;
  if (!c->n_nodes) return fail(c, "epv_set_tree must be called first");
  *bytes = halo_record_words(c->n_nodes, c->halo_k)*sizeof(unsigned long long);
  return 0;
}

int epv_halo_capacity(epv_ctx *c, uint64_t *capacity, uint64_t *needed) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  unsigned hmax = 0;
  EPV_CUDA(c, cudaMemcpyAsync(c->h_mail, c->d_hmax, sizeof(unsigned), cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  hmax = *reinterpret_cast<unsigned *>(c->h_mail);
  if (capacity) *capacity = (uint64_t)c->halo_k;
  // regrow while there is still a factor of two of headroom
  if (needed) *needed = std::max<uint64_t>((uint64_t)c->halo_k, 2*(uint64_t)hmax > (uint64_t)c->halo_k ? 4*(uint64_t)hmax + 16 : 0);
  return 0;
}

int epv_halo_need_async(epv_ctx *c, void *dst_device, void *stream) {
  if (!c) return 1;
  if (!c->n_sites) return fail(c, "no paths loaded");
  if (!dst_device) return fail(c, "epv_halo_need_async: dst_device is required");
  cudaSetDevice(c->device);
  cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : c->side_stream;
  EPV_CUDA(c, cudaMemsetAsync(dst_device, 0, sizeof(uint64_t), st));
  EPV_CUDA(c, cudaMemcpyAsync(dst_device, c->d_hmax, sizeof(unsigned), cudaMemcpyDeviceToDevice, st));   // low word
  return 0;
}

int epv_halo_set_capacity(epv_ctx *c, uint64_t capacity) {
  if (!c) return 1;
  if (!c->n_sites) return fail(c, "no paths loaded");
  if (capacity > 65000) return fail(c, "halo rows hold at most 65000 jumps");
  cudaSetDevice(c->device);
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  return set_halo_capacity(c, (int)capacity);
}

int epv_halo_pack(epv_ctx *c, int colour, void *rec_lo, void *rec_hi) {
  if (int r = halo_ready(c, colour)) return r;
  const int64_t own_lo = (int64_t)c->halo_lo, own_hi = (int64_t)(c->n_sites - c->halo_hi);
  int64_t site_lo = -1, site_hi = -1;
  if (c->halo_lo) {
    if (!rec_lo) return fail(c, "epv_halo_pack: this shard has a low neighbour, rec_lo is required");
    site_lo = own_lo + colour;
  }
  if (c->halo_hi) {
    if (!rec_hi) return fail(c, "epv_halo_pack: this shard has a high neighbour, rec_hi is required");
    site_hi = own_hi - 3 + (((colour - (own_hi - 3)) % 3) + 3) % 3;
  }
  if (site_lo < 0 && site_hi < 0) return 0;
  launch_halo_pack(halo_args(c), site_lo, static_cast<unsigned long long *>(rec_lo), site_hi,
                   static_cast<unsigned long long *>(rec_hi), c->stream);
  ++c->launches;
  EPV_CUDA(c, cudaGetLastError());
  return 0;
}

int epv_halo_unpack(epv_ctx *c, int colour, const void *rec_lo, const void *rec_hi) {
  if (int r = halo_ready(c, colour)) return r;
  const int64_t own_hi = (int64_t)(c->n_sites - c->halo_hi);
  int64_t site_lo = -1, site_hi = -1;
  if (c->halo_lo) {
    if (!rec_lo) return fail(c, "epv_halo_unpack: this shard has a low neighbour, rec_lo is required");
    site_lo = colour;
  }
  if (c->halo_hi) {
    if (!rec_hi) return fail(c, "epv_halo_unpack: this shard has a high neighbour, rec_hi is required");
    site_hi = own_hi + (((colour - own_hi) % 3) + 3) % 3;
  }
  if (site_lo < 0 && site_hi < 0) return 0;
  launch_halo_unpack(halo_args(c), site_lo, static_cast<const unsigned long long *>(rec_lo),
                     site_hi, static_cast<const unsigned long long *>(rec_hi), c->stream);
  ++c->launches;
  c->stats_valid = false;
  EPV_CUDA(c, cudaGetLastError());
  return 0;
}

}  // extern "C"
