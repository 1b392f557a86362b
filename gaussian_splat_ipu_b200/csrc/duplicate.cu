// duplicate.cu — offsets scan + K2: tile-overlap duplication.
//
// Replaces the reference's whole exchange phase (send/protocol/readInput/insert over the
// 4-neighbour lattice, codelets/splat/codelets.cpp:33-72,194-293,507-586 and
// src/splat/edge_builder.cpp:35-84): every drawn Gaussian is emitted once per tile of its
// converged rectangle as a 64-bit key (tile id << 32 | depth key) plus its index.
//
// Offsets come from a deterministic prefix sum over Gaussians in gid order (K1 leaves per-block
// sums, one small kernel scans them, K2 re-derives the in-block prefix), so that the stable
// radix sort that follows yields the canonical (z ascending, gid ascending) order per tile and
// the lists are bit-reproducible — an unordered atomic cursor would not give that.
#include "gsplat_internal.h"

namespace gsplat {

// Exclusive scan over 32-bit values for a block of THREADS threads; returns the exclusive prefix
// and writes the block total to *total (all threads).
template <int THREADS>
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t *warp_tot /* THREADS/32+1 */,
                                                    uint32_t &total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) warp_tot[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    uint32_t t = lane < THREADS / 32 ? warp_tot[lane] : 0, ti = t;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(0xffffffffu, ti, o);
      if (lane >= o) ti += u;
    }
    if (lane < THREADS / 32) warp_tot[lane] = ti - t;
    if (lane == THREADS / 32 - 1) warp_tot[THREADS / 32] = ti;
  }
  __syncthreads();
  total = warp_tot[THREADS / 32];
  const uint32_t r = warp_tot[warp] + inc - v;
  __syncthreads();   // warp_tot may be reused by the caller's next scan
  return r;
}

constexpr int SCAN_THREADS = 1024;
__global__ void __launch_bounds__(SCAN_THREADS)
scan_block_sums_kernel(uint32_t *__restrict__ sums, uint32_t nblocks, uint32_t *__restrict__ total) {
  __shared__ uint32_t wt[SCAN_THREADS / 32 + 1];
  uint32_t running = 0;
  for (uint32_t base = 0; base < nblocks; base += SCAN_THREADS) {
    const uint32_t i = base + threadIdx.x;
    const uint32_t v = i < nblocks ? sums[i] : 0;
    uint32_t tot;
    const uint32_t ex = block_excl_scan<SCAN_THREADS>(v, wt, tot);
    if (i < nblocks) sums[i] = running + ex;
    running += tot;
  }
  if (threadIdx.x == 0) *total = running;
}

void launch_scan_block_sums(uint32_t *block_sums, uint32_t nblocks, uint32_t *total, cudaStream_t st) {
  ++g_launches;
  scan_block_sums_kernel<<<1, SCAN_THREADS, 0, st>>>(block_sums, nblocks, total);
}

// K2.  One CTA covers the same 256 Gaussians as the K1 CTA of the same index.  The CTA's
// duplicates are a contiguous run [base, base + total) of the key array; the threads stride over
// that run and find the owning Gaussian by binary search in the in-block prefix, so writes are
// fully coalesced and a Gaussian covering hundreds of tiles does not serialise one thread.
__global__ void __launch_bounds__(K1_THREADS)
duplicate_kernel(Projected p, uint32_t n, const uint32_t *__restrict__ block_offsets, int tiles_across,
                 uint64_t *__restrict__ keys, uint32_t *__restrict__ vals, uint32_t capacity,
                 uint32_t *__restrict__ overflow) {
  __shared__ uint32_t wt[K1_THREADS / 32 + 1];
  __shared__ uint32_t s_prefix[K1_THREADS + 1];
  __shared__ uint32_t s_depth[K1_THREADS];
  __shared__ uint32_t s_cols[K1_THREADS];   // c0 | width << 16
  __shared__ uint32_t s_row0[K1_THREADS];
  const uint32_t tid = threadIdx.x;
  const uint32_t i = blockIdx.x * K1_THREADS + tid;
  const uint32_t t = i < n ? p.touched[i] : 0;
  uint32_t total;
  const uint32_t ex = block_excl_scan<K1_THREADS>(t, wt, total);
  if (total == 0) return;
  const uint32_t base = block_offsets[blockIdx.x];
  if (base + total > capacity || base + total < base) {   // never write past the buffer; host re-runs
    if (tid == 0) atomicExch(overflow, 1u);
    return;
  }
  s_prefix[tid] = ex;
  if (tid == 0) s_prefix[K1_THREADS] = total;
  if (t) {
    const float4 c = p.c[i];
    const uint32_t cols = __float_as_uint(c.z), rows = __float_as_uint(c.w);
    const uint32_t c0 = cols & 0xffffu, c1 = cols >> 16;
    // depth key: drawn Gaussians have clip z < 0 (codelets.cpp:493); ascending z (:314) ==
    // descending magnitude == ascending ~bits(z)
    s_depth[tid] = ~__float_as_uint(c.y);
    s_cols[tid] = c0 | ((c1 - c0 + 1) << 16);
    s_row0[tid] = rows & 0xffffu;
  }
  __syncthreads();
  for (uint32_t j = tid; j < total; j += K1_THREADS) {
    // largest g with s_prefix[g] <= j  (zero-count entries share their successor's prefix and are skipped)
    uint32_t lo = 0, hi = K1_THREADS;
#pragma unroll
    for (int s = 0; s < 8; ++s) {
      const uint32_t mid = (lo + hi) >> 1;
      if (s_prefix[mid] <= j) lo = mid; else hi = mid;
    }
    const uint32_t local = j - s_prefix[lo];
    const uint32_t cols = s_cols[lo];
    const uint32_t w = cols >> 16;
    const uint32_t ty = s_row0[lo] + local / w, tx = (cols & 0xffffu) + local % w;
    const uint32_t tile = ty * (uint32_t)tiles_across + tx;
    keys[base + j] = ((uint64_t)tile << 32) | s_depth[lo];
    vals[base + j] = blockIdx.x * K1_THREADS + lo;
  }
}

void launch_duplicate(const Projected &p, uint32_t n, const uint32_t *block_offsets, int tiles_across,
                      uint64_t *keys, uint32_t *vals, uint32_t capacity, uint32_t *overflow, cudaStream_t st) {
  if (!n) return;
  ++g_launches;
  duplicate_kernel<<<ceil_div(n, K1_THREADS), K1_THREADS, 0, st>>>(p, n, block_offsets, tiles_across, keys, vals,
                                                                  capacity, overflow);
}

}  // namespace gsplat
