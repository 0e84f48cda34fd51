// api_noalloc.cu -- ONE entry point that allocates nothing: the J/D sufficient-statistics pass
// (epv_suffstats, api_sampler.cu) with its scratch -- the integer accumulators the kernels add into
// before the fold -- in a file-scope device array instead of the context's cudaMalloc'ed buffer.
//
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
// that forbid it).  It has NO counterpart in the reference.
// No existing code path is changed: the kernels are the ones epv_suffstats launches, they
// only receive another scratch pointer.
//
// The array: EPV_STATIC_STATS_WORDS 8-byte words = 40 320 bytes, enough for trees of up to
// EPV_STATIC_MAX_NODES = 64 nodes (80 words per branch); every GPU that loads the library holds it until
// the process exits.  ONE array per GPU, shared by every context and every call on that GPU: it keeps the
// last call's contents (each call clears it first, on the context's stream), and two calls that overlap
// on the same GPU -- two contexts on different streams, two host threads -- RACE on it.  Serialise them.
//
// In host code the array's NAME is a host-side stand-in, not a device address (a kernel given it would
// fault, or silently use host memory where the GPU can reach it).  The device address is taken with
// cudaGetSymbolAddress in THIS file, once per device, with that device current, and passed to the kernels.
#include <mutex>

#include "../../include/epievo_b200.h"
#include "context.h"

using namespace epv;

namespace {

constexpr int STATIC_MAX_NODES = 64;
constexpr size_t STATIC_STATS_WORDS = (size_t)(STATIC_MAX_NODES - 1)*80;      // 80 = accumulator words per branch

static __device__ __align__(256) unsigned long long g_stats_scratch[STATIC_STATS_WORDS];

constexpr int MAX_DEVICES = 64;
std::mutex g_mutex;
unsigned long long *g_scratch_addr[MAX_DEVICES] = {};      // device address of g_stats_scratch, per device

// Device address of the array on `device` (which must be current); resolved once per device.
int scratch_address(epv_ctx *c, int device, unsigned long long **out) {
  if (device < 0 || device >= MAX_DEVICES) return fail(c, "epv_suffstats_static: device ordinal out of range");
  std::lock_guard<std::mutex> lock(g_mutex);
  if (!g_scratch_addr[device]) {
    unsigned long long *p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void **)&p, g_stats_scratch);
    if (e != cudaSuccess || !p)
      return fail(c, std::string("epv_suffstats_static: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
    g_scratch_addr[device] = p;
  }
  *out = g_scratch_addr[device];
  return 0;
}

}  // namespace

extern "C" {

int epv_static_limits(uint64_t *max_nodes, uint64_t *scratch_bytes) {
  if (max_nodes) *max_nodes = (uint64_t)STATIC_MAX_NODES;
  if (scratch_bytes) *scratch_bytes = (uint64_t)(STATIC_STATS_WORDS*sizeof(unsigned long long));
  return 0;
}

int epv_suffstats_static(epv_ctx *c, double *J, double *D) {
  if (!c) return 1;
  if (!J || !D) return fail(c, "epv_suffstats_static: J and D must not be NULL");
  if (!c->n_sites) return fail(c, "no paths loaded");
  if (c->n_nodes > STATIC_MAX_NODES ||
      suffstats_scratch_bytes(c->n_nodes, c->sm_count) > STATIC_STATS_WORDS*sizeof(unsigned long long))
    return fail(c, "epv_suffstats_static: the tree has " + std::to_string(c->n_nodes) + " nodes, the static "
                   "scratch holds at most " + std::to_string(STATIC_MAX_NODES) + " (use epv_suffstats)");
  EPV_CUDA(c, cudaSetDevice(c->device));
  unsigned long long *scratch = nullptr;
  if (int r = scratch_address(c, c->device, &scratch)) return r;
  // owned sites that have both neighbours globally (what epv_suffstats counts)
  uint64_t lo = c->upd_lo(), hi = c->upd_hi();
  const int64_t g0 = c->global0();
  if (g0 + (int64_t)lo < 1) lo = (uint64_t)(1 - g0);
  if (g0 + (int64_t)hi > (int64_t)c->n_global - 1) hi = (uint64_t)((int64_t)c->n_global - 1 - g0);
  // clear = true: the array still holds the previous call's sums
  c->launches += launch_suffstats(c->view(c->cur), c->d_state, c->W, c->tree(), c->n_sites, lo, hi, c->d_J, c->d_D,
                                  scratch, c->stream, c->sm_count, true, true);
  c->stats_valid = true;
  const size_t nb = (size_t)c->n_nodes*8*sizeof(double);
  EPV_CUDA(c, cudaMemcpyAsync(J, c->d_J, nb, cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaMemcpyAsync(D, c->d_D, nb, cudaMemcpyDeviceToHost, c->stream));
  EPV_CUDA(c, cudaStreamSynchronize(c->stream));
  EPV_CUDA(c, cudaGetLastError());
  return 0;
}

}  // extern "C"
