// radix_sort.cu — K3: hand-written device-wide stable LSD radix sort of (u64 key, u32 value)
// pairs, 8-bit digits.  Replaces the reference's per-tile quicksort by clip z
// (codelets/splat/codelets.cpp:303-356) together with the routing-by-tile the lattice did:
// one global sort of (tile id | depth) keys gives every tile its z-ordered list.
//
// Round-1 structure, three launches per digit: per-CTA digit histogram -> per-digit scan over
// CTAs -> ranked scatter.  No spin-waiting between CTAs.  HBM-bound: 8 B (hist) + 12 B + 12 B
// (scatter) per element per pass.  The element count is read on the device (*d_n) so the host
// never has to know D to launch the sort.
#include "gsplat_internal.h"

namespace gsplat {

__device__ __forceinline__ uint32_t lanemask_lt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

__device__ __forceinline__ uint32_t num_elems(const uint32_t *d_n, uint32_t cap) {
  const uint32_t n = *d_n;
  return n < cap ? n : cap;
}

// ---- pass kernel 1: digit histogram of one tile of RS_TILE keys -----------------------------
__global__ void __launch_bounds__(RS_THREADS)
rs_hist_kernel(const uint64_t *__restrict__ keys, const uint32_t *__restrict__ d_n, uint32_t cap, int shift,
               uint32_t *__restrict__ block_hist, uint32_t stride) {
  __shared__ uint32_t warp_hist[RS_THREADS / 32][256];
  const uint32_t n = num_elems(d_n, cap);
  const uint32_t b = blockIdx.x;
  if ((uint64_t)b * RS_TILE >= n) return;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int k = tid; k < (RS_THREADS / 32) * 256; k += RS_THREADS) (&warp_hist[0][0])[k] = 0;
  __syncthreads();
  const uint32_t seg = b * RS_TILE + warp * (32 * RS_ITEMS);
#pragma unroll 4
  for (int k = 0; k < RS_ITEMS; ++k) {
    const uint32_t idx = seg + k * 32 + lane;
    const bool valid = idx < n;
    const uint32_t d = valid ? (uint32_t)(keys[idx] >> shift) & 255u : 256u;
    const uint32_t m = __match_any_sync(0xffffffffu, d);
    if (valid && (m & lanemask_lt()) == 0) warp_hist[warp][d] += __popc(m);   // group leader
    __syncwarp();
  }
  __syncthreads();
  uint32_t c = 0;
#pragma unroll
  for (int w = 0; w < RS_THREADS / 32; ++w) c += warp_hist[w][tid];
  block_hist[(size_t)tid * stride + b] = c;
}

// ---- pass kernel 2: exclusive scan of each digit's row over CTAs; digit totals --------------
__global__ void __launch_bounds__(256)
rs_scan_kernel(uint32_t *__restrict__ block_hist, uint32_t stride, const uint32_t *__restrict__ d_n, uint32_t cap,
               uint32_t *__restrict__ digit_totals) {
  __shared__ uint32_t wt[256 / 32 + 1];
  const uint32_t n = num_elems(d_n, cap);
  const uint32_t nb = (uint32_t)(((uint64_t)n + RS_TILE - 1) / RS_TILE);
  uint32_t *row = block_hist + (size_t)blockIdx.x * stride;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint32_t running = 0;
  for (uint32_t base = 0; base < nb; base += 256) {
    const uint32_t i = base + tid;
    const uint32_t v = i < nb ? row[i] : 0;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) wt[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      uint32_t t = lane < 8 ? wt[lane] : 0, ti = t;
#pragma unroll
      for (int o = 1; o < 8; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, ti, o);
        if (lane >= o) ti += u;
      }
      if (lane < 8) wt[lane] = ti - t;
      if (lane == 7) wt[8] = ti;
    }
    __syncthreads();
    if (i < nb) row[i] = running + wt[warp] + inc - v;
    running += wt[8];
    __syncthreads();
  }
  if (tid == 0) digit_totals[blockIdx.x] = running;
}

// ---- pass kernel 3: stable ranked scatter ---------------------------------------------------
// Key order inside a tile: warp w owns the contiguous segment [w*512, (w+1)*512), item k / lane l
// is element k*32 + l of it, so (warp, k, lane) order == memory order and ranks are stable.
struct RsSmem {
  uint64_t keys[RS_TILE];
  uint32_t vals[RS_TILE];
  uint32_t warp_hist[RS_THREADS / 32][256];
  uint32_t tile_start[256];
  uint32_t out_base[256];
  uint32_t wt[RS_THREADS / 32 + 1];
};

__global__ void __launch_bounds__(RS_THREADS)
rs_scatter_kernel(const uint64_t *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
                  uint64_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out,
                  const uint32_t *__restrict__ d_n, uint32_t cap, int shift,
                  const uint32_t *__restrict__ block_hist, uint32_t stride,
                  const uint32_t *__restrict__ digit_totals) {
  extern __shared__ __align__(16) unsigned char rs_smem_raw[];
  RsSmem &sm = *reinterpret_cast<RsSmem *>(rs_smem_raw);
  const uint32_t n = num_elems(d_n, cap);
  const uint32_t b = blockIdx.x;
  if ((uint64_t)b * RS_TILE >= n) return;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t tile_n = min((uint32_t)RS_TILE, n - b * RS_TILE);

  for (int k = tid; k < (RS_THREADS / 32) * 256; k += RS_THREADS) (&sm.warp_hist[0][0])[k] = 0;
  __syncthreads();

  uint64_t key[RS_ITEMS];
  uint32_t rank[RS_ITEMS];
  const uint32_t seg = b * RS_TILE + warp * (32 * RS_ITEMS);
#pragma unroll
  for (int k = 0; k < RS_ITEMS; ++k) {
    const uint32_t idx = seg + k * 32 + lane;
    key[k] = idx < n ? keys_in[idx] : ~0ull;
  }
  const uint32_t lt = lanemask_lt();
#pragma unroll
  for (int k = 0; k < RS_ITEMS; ++k) {
    const bool valid = seg + k * 32 + lane < n;
    const uint32_t d = valid ? (uint32_t)(key[k] >> shift) & 255u : 256u;
    const uint32_t m = __match_any_sync(0xffffffffu, d);
    const int leader = __ffs(m) - 1;
    uint32_t old = 0;
    if (valid && (int)lane == leader) {
      old = sm.warp_hist[warp][d];
      sm.warp_hist[warp][d] = old + __popc(m);
    }
    old = __shfl_sync(0xffffffffu, old, leader);
    rank[k] = old + __popc(m & lt);
    __syncwarp();
  }
  __syncthreads();

  // digit `tid`: exclusive scan over warps, tile count, then exclusive scan over digits
  uint32_t cnt = 0;
#pragma unroll
  for (int w = 0; w < RS_THREADS / 32; ++w) {
    const uint32_t t = sm.warp_hist[w][tid];
    sm.warp_hist[w][tid] = cnt;
    cnt += t;
  }
  {
    uint32_t inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) sm.wt[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      uint32_t t = lane < 8 ? sm.wt[lane] : 0, ti = t;
#pragma unroll
      for (int o = 1; o < 8; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, ti, o);
        if (lane >= o) ti += u;
      }
      if (lane < 8) sm.wt[lane] = ti - t;
    }
    __syncthreads();
    const uint32_t tstart = sm.wt[warp] + inc - cnt;
    sm.tile_start[tid] = tstart;
    __syncthreads();
    // global base of digit `tid`: exclusive scan of the digit totals + this CTA's scanned row entry
    const uint32_t dt = digit_totals[tid];
    uint32_t dinc = dt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, dinc, o);
      if (lane >= o) dinc += t;
    }
    if (lane == 31) sm.wt[warp] = dinc;
    __syncthreads();
    if (warp == 0) {
      uint32_t t = lane < 8 ? sm.wt[lane] : 0, ti = t;
#pragma unroll
      for (int o = 1; o < 8; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, ti, o);
        if (lane >= o) ti += u;
      }
      if (lane < 8) sm.wt[lane] = ti - t;
    }
    __syncthreads();
    const uint32_t dbase = sm.wt[warp] + dinc - dt;
    sm.out_base[tid] = dbase + block_hist[(size_t)tid * stride + b] - tstart;
  }
  __syncthreads();

  // reorder through shared memory so the global writes are runs of equal digits
#pragma unroll
  for (int k = 0; k < RS_ITEMS; ++k) {
    const uint32_t idx = seg + k * 32 + lane;
    if (idx < n) {
      const uint32_t d = (uint32_t)(key[k] >> shift) & 255u;
      const uint32_t pos = sm.tile_start[d] + sm.warp_hist[warp][d] + rank[k];
      sm.keys[pos] = key[k];
      sm.vals[pos] = vals_in[idx];
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < RS_ITEMS; ++k) {
    const uint32_t pos = k * RS_THREADS + tid;
    if (pos < tile_n) {
      const uint64_t kk = sm.keys[pos];
      const uint32_t d = (uint32_t)(kk >> shift) & 255u;
      const uint32_t dst = sm.out_base[d] + pos;
      keys_out[dst] = kk;
      vals_out[dst] = sm.vals[pos];
    }
  }
}

int launch_radix_sort(SortBuffers &sb, const uint32_t *d_n, uint32_t n_upper, int key_bits, cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(rs_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(RsSmem));
    attr_set = true;
  }
  if (n_upper > sb.capacity) n_upper = sb.capacity;
  const uint32_t nb = (uint32_t)(((uint64_t)n_upper + RS_TILE - 1) / RS_TILE);
  int cur = 0;
  if (nb == 0) return cur;
  const int passes = (key_bits + 7) / 8;
  for (int p = 0; p < passes; ++p) {
    const int shift = p * 8;
    ++g_launches;
    rs_hist_kernel<<<nb, RS_THREADS, 0, st>>>(sb.keys[cur], d_n, sb.capacity, shift, sb.block_hist, sb.max_blocks);
    ++g_launches;
    rs_scan_kernel<<<256, 256, 0, st>>>(sb.block_hist, sb.max_blocks, d_n, sb.capacity, sb.digit_base);
    ++g_launches;
    rs_scatter_kernel<<<nb, RS_THREADS, sizeof(RsSmem), st>>>(sb.keys[cur], sb.vals[cur], sb.keys[cur ^ 1],
                                                              sb.vals[cur ^ 1], d_n, sb.capacity, shift,
                                                              sb.block_hist, sb.max_blocks, sb.digit_base);
    cur ^= 1;
  }
  return cur;
}

// ---- K4: per-tile ranges over the sorted keys (counts = end - start) ---------
__global__ void __launch_bounds__(256)
tile_ranges_kernel(const uint64_t *__restrict__ keys, const uint32_t *__restrict__ d_n, uint32_t cap,
                   uint2 *__restrict__ ranges, uint32_t num_tiles) {
  const uint32_t n = num_elems(d_n, cap);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t t = (uint32_t)(keys[i] >> 32);
    if (t >= num_tiles) continue;
    if (i == 0 || (uint32_t)(keys[i - 1] >> 32) != t) ranges[t].x = i;
    if (i == n - 1 || (uint32_t)(keys[i + 1] >> 32) != t) ranges[t].y = i + 1;
  }
}

void launch_tile_ranges(const uint64_t *keys, const uint32_t *d_n, uint32_t n_upper, uint32_t cap, uint2 *ranges,
                        uint32_t num_tiles, cudaStream_t st) {
  cudaMemsetAsync(ranges, 0, sizeof(uint2) * num_tiles, st);
  if (!n_upper) return;
  const int blocks = min(ceil_div(n_upper, 256), 148 * 8);
  ++g_launches;
  tile_ranges_kernel<<<blocks, 256, 0, st>>>(keys, d_n, cap, ranges, num_tiles);
}

}  // namespace gsplat
