// kernels_compact.cu -- the compact host exchange format (2 bytes per row instead of 9):
//   row_code u16 [n_nodes * n_sites]   jump count of the row | init_state << 15 (root rows: count 0)
//   jumps    f64 [sum of counts]       in row order (node-major), as in the CSR format
// Upload: per-branch totals -> (host sizes the buffers) -> offsets by a single-pass chained
// scan + state bit sets, all on the device.  Download: the inverse, one thread per site.
#include "kernels.h"
#include "scan.cuh"

namespace epv {

namespace {

// totals[bi*3 + {0,1,2}] = jumps in the low-halo rows, the owned rows, the high-halo rows
__global__ void __launch_bounds__(256)
compact_totals_kernel(const uint16_t *code, uint64_t n_sites, uint64_t own_lo, uint64_t own_hi,
                      unsigned long long *totals) {
  const int bi = blockIdx.y;
  const uint16_t *row = code + (uint64_t)(bi + 1)*n_sites;
  unsigned long long t[3] = {0, 0, 0};
  for (uint64_t s = (uint64_t)blockIdx.x*blockDim.x + threadIdx.x; s < n_sites;
       s += (uint64_t)gridDim.x*blockDim.x) {
    const unsigned len = row[s] & 0x7fffu;
    t[s < own_lo ? 0 : s < own_hi ? 1 : 2] += len;
  }
  for (int k = 0; k < 3; ++k) {
    const unsigned long long v = warp_sum(t[k]);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(totals + bi*3 + k, v);
  }
}

// Owned rows -> CSR offsets (thread = 4 consecutive rows, as in csr_rebuild_kernel) and the
// state bit sets of those sites (64 nodes per pass over the branches).
__global__ void __launch_bounds__(RB_THREADS)
compact_import_kernel(const uint16_t *code, uint64_t n_sites, int n_nodes, uint64_t own_lo,
                      uint64_t n_own, uint32_t *off, uint64_t pitch, uint8_t *state, int W,
                      unsigned long long *desc, uint32_t n_tiles, uint32_t epoch, uint32_t *ticket,
                      uint32_t ticket_base) {
  extern __shared__ uint32_t s_len[];            // [n_branches][RB_THREADS]
  __shared__ uint32_t s_tile;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int n_br = n_nodes - 1;
  if (t == 0) s_tile = atomicAdd(ticket, 1u) - ticket_base;
  __syncthreads();
  const uint32_t tile = s_tile;
  const uint64_t R = 4*((uint64_t)tile*RB_THREADS + t);

  // state bits: node v of site s is bit 15 of code[v*n_sites + s]
  for (int word = 0; word*64 < n_nodes; ++word) {
    uint64_t sb[4] = {0, 0, 0, 0};
    for (int v = word*64; v < min(n_nodes, word*64 + 64); ++v)
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (R + k < n_own)
          sb[k] |= (uint64_t)(code[(uint64_t)v*n_sites + own_lo + R + k] >> 15) << (v & 63);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (R + k >= n_own) continue;
      uint8_t *p = state + (own_lo + R + k)*W;
      switch (W) {
      case 1: *p = (uint8_t)sb[k]; break;
      case 2: *reinterpret_cast<uint16_t *>(p) = (uint16_t)sb[k]; break;
      case 4: *reinterpret_cast<uint32_t *>(p) = (uint32_t)sb[k]; break;
      default: reinterpret_cast<uint64_t *>(p)[word] = sb[k];
      }
    }
  }

  for (int bi = 0; bi < n_br; ++bi) {
    uint32_t sum = 0;
    const uint16_t *row = code + (uint64_t)(bi + 1)*n_sites + own_lo;
#pragma unroll
    for (int k = 0; k < 4; ++k) if (R + k < n_own) sum += row[R + k] & 0x7fffu;
    s_len[bi*RB_THREADS + t] = sum;
  }
  __syncthreads();
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
    const uint32_t base = tile_base(desc + (uint64_t)bi*n_tiles, tile, epoch, tot, lane);
    uint32_t run = base + incl - mine;
#pragma unroll
    for (int k = 0; k < PER; ++k) { row[lane*PER + k] = run; run += v[k]; }
  }
  __syncthreads();
  if (R > n_own) return;
  for (int bi = 0; bi < n_br; ++bi) {
    const uint16_t *row = code + (uint64_t)(bi + 1)*n_sites + own_lo;
    uint32_t w = s_len[bi*RB_THREADS + t], out[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { out[k] = w; if (R + k < n_own) w += row[R + k] & 0x7fffu; }
    *reinterpret_cast<uint4 *>(off + (uint64_t)bi*pitch + R) = make_uint4(out[0], out[1], out[2], out[3]);
  }
}

__global__ void __launch_bounds__(256)
compact_export_kernel(const PathView pv, const uint8_t *state, int W, uint64_t n_sites, int n_nodes,
                      uint16_t *code, unsigned *err) {
  const uint64_t s = (uint64_t)blockIdx.x*blockDim.x + threadIdx.x;
  if (s >= n_sites) return;
  code[s] = (uint16_t)((uint32_t)state_bit(state, W, s, 0) << 15);
  for (int v = 1; v < n_nodes; ++v) {
    const int bi = v - 1;
    uint32_t len;
    if (s >= pv.own_lo && s < pv.own_hi) {
      const uint32_t *o = pv.off + (uint64_t)bi*pv.off_pitch + (s - pv.own_lo);
      len = o[1] - o[0];
    }
    else len = pv.hcnt[(s < pv.own_lo ? (int)s : HALO_W + (int)(s - pv.own_hi))*pv.n_br + bi];
    if (len > 0x7fffu) { atomicOr(err, ERR_BAD_INPUT); len = 0x7fffu; }
    code[(uint64_t)v*n_sites + s] = (uint16_t)(len | (uint32_t)state_bit(state, W, s, v) << 15);
  }
}

}  // namespace

void
launch_compact_totals(const uint16_t *code, uint64_t n_sites, int n_br, uint64_t own_lo,
                      uint64_t own_hi, unsigned long long *totals, cudaStream_t st) {
  cudaMemsetAsync(totals, 0, (size_t)n_br*3*sizeof(unsigned long long), st);
  const unsigned bx = (unsigned)((n_sites + 256*16 - 1)/(256*16));
  compact_totals_kernel<<<dim3(bx ? bx : 1, n_br), 256, 0, st>>>(code, n_sites, own_lo, own_hi, totals);
}

void
launch_compact_import(const uint16_t *code, uint64_t n_sites, int n_nodes, uint64_t own_lo,
                      uint64_t n_own, uint32_t *off, uint64_t pitch, uint8_t *state, int W,
                      unsigned long long *desc, uint32_t n_tiles, uint32_t epoch, uint32_t *ticket,
                      uint32_t ticket_base, cudaStream_t st) {
  const size_t smem = (size_t)(n_nodes - 1)*RB_THREADS*sizeof(uint32_t);
  cudaFuncSetAttribute(compact_import_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  compact_import_kernel<<<n_tiles, RB_THREADS, smem, st>>>(code, n_sites, n_nodes, own_lo, n_own, off,
                                                         pitch, state, W, desc, n_tiles, epoch, ticket,
                                                         ticket_base);
}

void
launch_compact_export(const PathView &pv, const uint8_t *state, int W, uint64_t n_sites, int n_nodes,
                      uint16_t *code, unsigned *err, cudaStream_t st) {
  compact_export_kernel<<<(unsigned)((n_sites + 255)/256), 256, 0, st>>>(pv, state, W, n_sites, n_nodes,
                                                                        code, err);
}

}  // namespace epv
