// duplicate.cu — K2: tile-overlap duplication.
//
// Replaces the reference's whole exchange phase (send/protocol/readInput/insert over the
// 4-neighbour lattice, codelets/splat/codelets.cpp:33-72,194-293,507-586 and
// src/splat/edge_builder.cpp:35-84): every drawn Gaussian is emitted once per tile of its
// converged rectangle as a (tile id, Gaussian index) pair.
//
// K2 walks the Gaussians in DEPTH order (the permutation sort A produced), so the emitted array is
// globally ordered by (clip z, gid) and the stable sort by tile id that follows (sort B) leaves
// every tile's list in the canonical (z ascending, gid ascending) order.  Offsets come from a
// single-pass chained scan with decoupled look-back — deterministic prefix sums, not an unordered
// atomic cursor — so the lists are bit-reproducible.  CTAs take their position in the chain from an
// atomic ticket, which guarantees that every predecessor a CTA waits on is already running.
#include "gsplat_internal.h"

namespace gsplat {

// Exclusive scan over 32-bit values for a block of THREADS threads; returns the exclusive prefix
// and the block total (all threads).
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

// One CTA = one link of the chain = K2_CHUNK consecutive positions of the depth order
// (K2_ITEMS per thread, striped so the loads coalesce and the rectangle gathers overlap).
//
// Depth slices.  K5 stops a pixel at the first Gaussian that would push its transmittance under
// 1e-4 (codelets.cpp:405-408); once every pixel of a tile has stopped, nothing further down the
// tile's list can change the image.  The frame is therefore binned and blended in a few slices of
// increasing depth (select.cu): one launch walks the depth-sorted Gaussians of one slice and, when
// CHECK_DONE, emits nothing for tiles whose `tile_done` flag K5 raised in an earlier slice.
// Every tile's list is then an exact PREFIX of the reference's list — complete unless the tile
// saturated — and the framebuffer is bit-identical to the one-slice run.  The per-tile COUNTS of
// the reference (codelets.cpp:501-504) still count every Gaussian: tile_count_kernel below.
constexpr uint32_t K2_SMALL = 12;   // rectangles up to this many tiles are written by their own thread
struct K2Smem {
  uint32_t prefix[K2_CHUNK];
  uint32_t wt[K2_THREADS / 32 + 1];
  uint32_t chain;            // this CTA's position in the chain, then its base offset
  uint32_t n_big;
  uint32_t big[K2_CHUNK];    // chunk entries with more than K2_SMALL tiles (usually none)
  uint2 big_rect[K2_CHUNK];
  uint32_t big_gauss[K2_CHUNK];
};

template <bool CHECK_DONE>
__device__ __forceinline__ bool tile_live(const uint8_t *__restrict__ tile_done, uint32_t tile) {
  if (!CHECK_DONE) return true;
  return __ldg(&tile_done[tile]) == 0;
}

// Writes the live tiles of one big rectangle, row-major, starting at `at`: one warp, 32 tiles per
// step, positions from a ballot prefix (so skipped tiles leave no gap).
template <bool CHECK_DONE>
__device__ __forceinline__ void emit_rect_warp(uint32_t *__restrict__ keys, uint32_t *__restrict__ vals, uint32_t at,
                                               uint2 rc, uint32_t g, int tiles_across,
                                               const uint8_t *__restrict__ tile_done, uint32_t lane) {
  const uint32_t c0 = rc.x & 0xffffu, w = (rc.x >> 16) - c0 + 1, r0 = rc.y & 0xffffu;
  const uint32_t total = w * ((rc.y >> 16) - r0 + 1);
  for (uint32_t q0 = 0; q0 < total; q0 += 32) {
    const uint32_t q = q0 + lane;
    uint32_t tile = 0;
    bool live = false;
    if (q < total) {
      tile = (r0 + q / w) * (uint32_t)tiles_across + c0 + q % w;
      live = tile_live<CHECK_DONE>(tile_done, tile);
    }
    const uint32_t m = __ballot_sync(0xffffffffu, live);
    if (live) {
      const uint32_t o = at + __popc(m & ((1u << lane) - 1u));
      keys[o] = tile;
      vals[o] = g;
    }
    at += __popc(m);
  }
}

template <bool CHECK_DONE>
__global__ void __launch_bounds__(K2_THREADS)
duplicate_kernel(Projected p, const uint32_t *__restrict__ sorted0, const uint32_t *__restrict__ sorted1,
                 const uint32_t *__restrict__ key_range, uint32_t sort_bpp, uint32_t sort_passes,
                 const uint32_t *__restrict__ d_count,
                 uint32_t count_cap, int tiles_across, const uint8_t *__restrict__ tile_done,
                 const uint32_t *__restrict__ live_sat, uint32_t *__restrict__ keys, uint32_t *__restrict__ vals, uint32_t capacity,
                 unsigned long long *__restrict__ lookback, uint32_t *__restrict__ ticket,
                 uint32_t *__restrict__ total_out, uint32_t *__restrict__ overflow) {
  extern __shared__ __align__(16) unsigned char k2_smem_raw[];
  K2Smem &sm = *reinterpret_cast<K2Smem *>(k2_smem_raw);
  const uint32_t tid = threadIdx.x;
  // the slice's Gaussian count is only known on the device: the grid covers the capacity and the
  // CTAs beyond the last chunk leave at once (before taking a chain ticket)
  const uint32_t end = min(*d_count, count_cap);
  const uint32_t nblk = (end + K2_CHUNK - 1) / K2_CHUNK;
  if (blockIdx.x >= nblk) return;
  if (tid == 0) { sm.chain = atomicAdd(ticket, 1u); sm.n_big = 0; }
  __syncthreads();
  const uint32_t b = sm.chain;
  // where the depth sort left its result (it skips the passes above the key range on the device)
  const uint32_t *__restrict__ sorted_gauss =
      key_range && rs_result_buffer(key_range, sort_bpp, sort_passes) ? sorted1 : sorted0;
  uint32_t g[K2_ITEMS];
  uint2 rc[K2_ITEMS];
  uint32_t cnt[K2_ITEMS];
  bool live[K2_ITEMS];
#pragma unroll
  for (int i = 0; i < K2_ITEMS; ++i) {
    const uint32_t j = b * K2_CHUNK + i * K2_THREADS + tid;   // position in the slice's depth order
    live[i] = j < end;
    g[i] = live[i] ? sorted_gauss[j] : 0u;
  }
#pragma unroll
  for (int i = 0; i < K2_ITEMS; ++i) {
    rc[i] = live[i] ? __ldg(&p.rect[g[i]]) : make_uint2(0u, 0u);
    const uint32_t c0 = rc[i].x & 0xffffu, c1 = rc[i].x >> 16, r0 = rc[i].y & 0xffffu, r1 = rc[i].y >> 16;
    uint32_t c = 0;
    if (live[i] && r1 >= r0) {
      if (!CHECK_DONE) c = (c1 - c0 + 1) * (r1 - r0 + 1);
      else c = live_tiles_in_rect(live_sat, tiles_across, c0, c1, r0, r1);   // tiles still being blended
    }
    cnt[i] = c;
    sm.prefix[i * K2_THREADS + tid] = c;
  }
  __syncthreads();
  // exclusive scan of the chunk in position order: thread tid owns entries [tid*ITEMS, (tid+1)*ITEMS)
  uint32_t loc[K2_ITEMS], sum = 0;
#pragma unroll
  for (int i = 0; i < K2_ITEMS; ++i) { loc[i] = sm.prefix[tid * K2_ITEMS + i]; sum += loc[i]; }
  uint32_t total;
  uint32_t ex = block_excl_scan<K2_THREADS>(sum, sm.wt, total);
#pragma unroll
  for (int i = 0; i < K2_ITEMS; ++i) { sm.prefix[tid * K2_ITEMS + i] = ex; ex += loc[i]; }

  // ---- chained scan: publish the aggregate, look back for the exclusive prefix ------------------
  if (tid < 32) {
    const unsigned long long excl = lb_exclusive_prefix(lookback, b, total, tid);
    if (tid == 0) {
      const unsigned long long incl = excl + total;
      if (b == nblk - 1) *total_out = incl > 0xffffffffull ? 0xffffffffu : (uint32_t)incl;
      if (incl > capacity) atomicExch(overflow, 1u);   // never write past the buffer; the host re-runs
      sm.chain = incl > capacity ? 0xffffffffu : (uint32_t)excl;
    }
  }
  __syncthreads();
  const uint32_t base = sm.chain;
  if (total == 0 || base == 0xffffffffu) return;

  // The CTA's duplicates are the contiguous run [base, base + total).  Neighbouring threads own
  // neighbouring entries, so a warp's stores fall into one short contiguous span.  Small
  // rectangles (the common case: 3.4 tiles on average) are written by their own thread; the rare
  // large ones are queued and written warp by warp so they do not serialise one thread.
#pragma unroll
  for (int i = 0; i < K2_ITEMS; ++i) {
    if (cnt[i] == 0) continue;
    const uint32_t e = i * K2_THREADS + tid;
    const uint32_t c0 = rc[i].x & 0xffffu, c1 = rc[i].x >> 16;
    const uint32_t area = (c1 - c0 + 1) * ((rc[i].y >> 16) - (rc[i].y & 0xffffu) + 1);
    if (area <= K2_SMALL) {   // row-major walk of the rectangle without divisions
      uint32_t at = base + sm.prefix[e], tx = c0, row = (rc[i].y & 0xffffu) * (uint32_t)tiles_across;
      for (uint32_t q = 0; q < area; ++q) {
        if (tile_live<CHECK_DONE>(tile_done, row + tx)) {
          keys[at] = row + tx;
          vals[at] = g[i];
          ++at;
        }
        if (tx == c1) { tx = c0; row += (uint32_t)tiles_across; } else { ++tx; }
      }
    } else {
      const uint32_t slot = atomicAdd(&sm.n_big, 1u);
      sm.big[slot] = e; sm.big_rect[slot] = rc[i]; sm.big_gauss[slot] = g[i];
    }
  }
  __syncthreads();
  const uint32_t n_big = sm.n_big;
  for (uint32_t s = tid >> 5; s < n_big; s += K2_THREADS / 32)
    emit_rect_warp<CHECK_DONE>(keys, vals, base + sm.prefix[sm.big[s]], sm.big_rect[s], sm.big_gauss[s], tiles_across,
                               tile_done, tid & 31);
}

void launch_duplicate(const Projected &p, const uint32_t *const sorted_gauss[2], const uint32_t *key_range,
                      uint32_t sort_bits_per_pass, uint32_t sort_passes, const uint32_t *d_count, uint32_t count_cap,
                      int tiles_across, const uint8_t *tile_done, const uint32_t *live_sat, uint32_t *keys, uint32_t *vals,
                      uint32_t capacity, unsigned long long *lookback, uint32_t *ticket, uint32_t *total,
                      uint32_t *overflow, cudaStream_t st) {
  if (!count_cap) return;
  ++g_launches;
  const int grid = ceil_div(count_cap, K2_CHUNK);
  if (tile_done)
    duplicate_kernel<true><<<grid, K2_THREADS, sizeof(K2Smem), st>>>(p, sorted_gauss[0], sorted_gauss[1], key_range,
                                                                     sort_bits_per_pass, sort_passes, d_count, count_cap, tiles_across,
                                                                     tile_done, live_sat, keys, vals, capacity, lookback,
                                                                     ticket, total, overflow);
  else
    duplicate_kernel<false><<<grid, K2_THREADS, sizeof(K2Smem), st>>>(p, sorted_gauss[0], sorted_gauss[1], key_range,
                                                                      sort_bits_per_pass, sort_passes, d_count, count_cap, tiles_across,
                                                                      nullptr, nullptr, keys, vals, capacity, lookback,
                                                                      ticket, total, overflow);
}

// ---- a11: per-tile counts of the reference (codelets.cpp:501-504; getIPUHistogram) ---------------
// counts[tile] = number of drawn Gaussians whose rectangle covers the tile = the length of the
// reference's render list, whether or not the depth slices ever emit that far.  Each rectangle
// adds +1/-1 at its four corners of a (rows+1) x (cols+1) difference grid (shared-memory copy per
// CTA, flushed with one atomic per touched cell); a 2-D prefix sum turns the grid into counts.
constexpr int TC_THREADS = 1024;
constexpr int TC_BATCH = 4;
__global__ void __launch_bounds__(TC_THREADS)
tile_count_kernel(Projected p, uint32_t n, int tiles_across, int tiles_down, int *__restrict__ diff) {
  extern __shared__ int sdiff[];
  const int gw = tiles_across + 1, cells = gw * (tiles_down + 1);
  for (int k = threadIdx.x; k < cells; k += blockDim.x) sdiff[k] = 0;
  __syncthreads();
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t i0 = blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += stride * TC_BATCH) {
    uint2 rc[TC_BATCH];
#pragma unroll
    for (int k = 0; k < TC_BATCH; ++k) {   // TC_BATCH independent loads in flight per thread
      const uint32_t i = i0 + k * stride;
      rc[k] = i < n ? p.rect[i] : make_uint2(RECT_EMPTY, RECT_EMPTY);
    }
#pragma unroll
    for (int k = 0; k < TC_BATCH; ++k) {
      const int c0 = rc[k].x & 0xffffu, c1 = rc[k].x >> 16, r0 = rc[k].y & 0xffffu, r1 = rc[k].y >> 16;
      if (c1 < c0 || r1 < r0) continue;   // not drawn (RECT_EMPTY), or outside this context's band
      atomicAdd(&sdiff[r0 * gw + c0], 1);
      atomicAdd(&sdiff[r0 * gw + c1 + 1], -1);
      atomicAdd(&sdiff[(r1 + 1) * gw + c0], -1);
      atomicAdd(&sdiff[(r1 + 1) * gw + c1 + 1], 1);
    }
  }
  __syncthreads();
  for (int k = threadIdx.x; k < cells; k += blockDim.x)
    if (sdiff[k]) atomicAdd(&diff[k], sdiff[k]);
}

// single CTA: 2-D inclusive prefix of the difference grid -> counts; clears the grid for the next frame
__global__ void __launch_bounds__(1024)
tile_count_finish_kernel(int *__restrict__ diff, int tiles_across, int tiles_down, uint32_t *__restrict__ counts) {
  extern __shared__ int sgrid[];
  const int gw = tiles_across + 1, gh = tiles_down + 1, cells = gw * gh;
  for (int k = threadIdx.x; k < cells; k += blockDim.x) { sgrid[k] = diff[k]; diff[k] = 0; }
  __syncthreads();
  for (int r = threadIdx.x; r < gh; r += blockDim.x) {
    int run = 0;
    for (int c = 0; c < gw; ++c) { run += sgrid[r * gw + c]; sgrid[r * gw + c] = run; }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < gw; c += blockDim.x) {
    int run = 0;
    for (int r = 0; r < gh; ++r) { run += sgrid[r * gw + c]; sgrid[r * gw + c] = run; }
  }
  __syncthreads();
  for (int k = threadIdx.x; k < tiles_across * tiles_down; k += blockDim.x)
    counts[k] = (uint32_t)sgrid[(k / tiles_across) * gw + k % tiles_across];
}

size_t tile_count_smem(int tiles_across, int tiles_down);
// Summed-area table of the tiles that are still being blended: sat[r * (cols+1) + c] = live tiles
// in rows < r and columns < c.  Later slices ask "how many live tiles does this rectangle cover"
// with four reads (select.cu, K2) instead of walking the rectangle.  Single CTA.
__global__ void __launch_bounds__(1024)
live_sat_kernel(const uint8_t *__restrict__ tile_done, int tiles_across, int tiles_down, uint32_t *__restrict__ sat) {
  extern __shared__ int sgrid[];
  const int gw = tiles_across + 1, gh = tiles_down + 1, cells = gw * gh;
  for (int k = threadIdx.x; k < cells; k += blockDim.x) {
    const int r = k / gw, c = k % gw;
    sgrid[k] = (r > 0 && c > 0 && tile_done[(r - 1) * tiles_across + (c - 1)] == 0) ? 1 : 0;
  }
  __syncthreads();
  for (int r = threadIdx.x; r < gh; r += blockDim.x) {
    int run = 0;
    for (int c = 0; c < gw; ++c) { run += sgrid[r * gw + c]; sgrid[r * gw + c] = run; }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < gw; c += blockDim.x) {
    int run = 0;
    for (int r = 0; r < gh; ++r) { run += sgrid[r * gw + c]; sgrid[r * gw + c] = run; }
  }
  __syncthreads();
  for (int k = threadIdx.x; k < cells; k += blockDim.x) sat[k] = (uint32_t)sgrid[k];
}

void launch_live_sat(const uint8_t *tile_done, int tiles_across, int tiles_down, uint32_t *sat, cudaStream_t st) {
  ++g_launches;
  live_sat_kernel<<<1, 1024, tile_count_smem(tiles_across, tiles_down), st>>>(tile_done, tiles_across, tiles_down, sat);
}

// Dynamic shared memory opt-in of K2 and the tile-grid kernels on the CURRENT device (per device:
// gsplat_create runs this once for each).
cudaError_t setup_duplicate_kernels() {
  cudaError_t e = cudaSuccess;
  auto acc = [&](cudaError_t r) { if (e == cudaSuccess) e = r; };
  acc(cudaFuncSetAttribute(duplicate_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(K2Smem)));
  acc(cudaFuncSetAttribute(duplicate_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(K2Smem)));
  acc(cudaFuncSetAttribute(live_sat_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TILE_COUNT_MAX_SMEM));
  acc(cudaFuncSetAttribute(tile_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TILE_COUNT_MAX_SMEM));
  acc(cudaFuncSetAttribute(tile_count_finish_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TILE_COUNT_MAX_SMEM));
  return e;
}

size_t tile_count_smem(int tiles_across, int tiles_down) {
  return sizeof(int) * (size_t)(tiles_across + 1) * (tiles_down + 1);
}

void launch_tile_counts(const Projected &p, uint32_t n, int tiles_across, int tiles_down, int *diff, uint32_t *counts,
                        cudaStream_t st) {
  const size_t smem = tile_count_smem(tiles_across, tiles_down);
  g_launches += n ? 2 : 1;
  if (n) tile_count_kernel<<<148 * 2, TC_THREADS, smem, st>>>(p, n, tiles_across, tiles_down, diff);
  tile_count_finish_kernel<<<1, 1024, smem, st>>>(diff, tiles_across, tiles_down, counts);
}

}  // namespace gsplat
