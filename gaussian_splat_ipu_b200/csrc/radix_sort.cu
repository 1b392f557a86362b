// radix_sort.cu — K3: hand-written device-wide stable LSD radix sort of (u32 key, u32 value)
// pairs, and K4 (per-tile ranges).
//
// The reference orders each tile's list with a per-tile quicksort by clip z
// (codelets/splat/codelets.cpp:303-356) after its lattice has routed Gaussians to tiles.  Here the
// same order — (tile, z ascending, gid ascending) — comes from TWO uses of this one sort:
//   sort A  the N Gaussians by their 32-bit depth key (4 digits of 8 bits), values = Gaussian index
//   sort B  the D (tile, Gaussian) duplicates, emitted by K2 in depth order, by tile id only
//           (ceil(log2(tiles)) bits: 2 digits of 6 bits at 1080p, of 7 bits at 4K)
// Both are stable, so sort B keeps the depth order inside every tile: the result is exactly what a
// single stable sort of 64-bit (tile << 32 | depth) keys gives, at (64·N + 40·D) instead of
// 152·D bytes of HBM traffic.
//
// Structure per digit, three launches: per-tile digit histogram -> per-digit exclusive scan over
// tiles -> ranked scatter.  "Tile" here = RS_TILE consecutive elements; CTAs walk tiles with a grid
// stride and the element count is read on the device (*d_n), so the host never needs to know N or
// D to launch the sort (no mid-frame synchronisation).  HBM-bound: 4 B (hist) + 8 B + 8 B (scatter)
// per element per digit.
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

constexpr int RS_WARPS = RS_THREADS / 32;

// ---- digit histogram of every tile -----------------------------------------------------------
__global__ void __launch_bounds__(RS_THREADS)
rs_hist_kernel(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ d_n, uint32_t cap, int shift,
               uint32_t mask, uint32_t *__restrict__ block_hist, uint32_t stride) {
  __shared__ uint32_t warp_hist[RS_WARPS][256];
  const uint32_t n = num_elems(d_n, cap);
  const uint32_t ntiles = (uint32_t)(((uint64_t)n + RS_TILE - 1) / RS_TILE);
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (uint32_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
    for (int k = tid; k < RS_WARPS * 256; k += RS_THREADS) (&warp_hist[0][0])[k] = 0;
    __syncthreads();
    const uint32_t seg = t * RS_TILE + warp * (32 * RS_ITEMS);
#pragma unroll 4
    for (int k = 0; k < RS_ITEMS; ++k) {
      const uint32_t idx = seg + k * 32 + lane;
      const bool valid = idx < n;
      const uint32_t d = valid ? (keys[idx] >> shift) & mask : 256u;
      const uint32_t m = __match_any_sync(0xffffffffu, d);
      if (valid && (m & lanemask_lt()) == 0) warp_hist[warp][d] += __popc(m);   // group leader
      __syncwarp();
    }
    __syncthreads();
    if (tid <= mask) {
      uint32_t c = 0;
#pragma unroll
      for (int w = 0; w < RS_WARPS; ++w) c += warp_hist[w][tid];
      block_hist[(size_t)tid * stride + t] = c;
    }
    __syncthreads();
  }
}

// ---- exclusive scan of each digit's row over tiles; digit totals -------------------------------
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

// ---- stable ranked scatter ---------------------------------------------------------------------
// Key order inside a tile: warp w owns the contiguous segment [w*512, (w+1)*512), item k / lane l
// is element k*32 + l of it, so (warp, k, lane) order == memory order and ranks are stable.
struct RsSmem {
  uint32_t keys[RS_TILE];
  uint32_t vals[RS_TILE];
  uint32_t warp_hist[RS_WARPS][256];
  uint32_t tile_start[256];
  uint32_t out_base[256];
  uint32_t wt[RS_WARPS + 1];
};

// exclusive scan of one value per thread over the 256-thread CTA (result for every thread)
__device__ __forceinline__ uint32_t cta_excl_scan256(uint32_t v, uint32_t *wt, uint32_t lane, uint32_t warp) {
  uint32_t inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) wt[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    uint32_t t = lane < RS_WARPS ? wt[lane] : 0, ti = t;
#pragma unroll
    for (int o = 1; o < RS_WARPS; o <<= 1) {
      const uint32_t u = __shfl_up_sync(0xffffffffu, ti, o);
      if (lane >= o) ti += u;
    }
    if (lane < RS_WARPS) wt[lane] = ti - t;
  }
  __syncthreads();
  const uint32_t r = wt[warp] + inc - v;
  __syncthreads();
  return r;
}

__global__ void __launch_bounds__(RS_THREADS)
rs_scatter_kernel(const uint32_t *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
                  uint32_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out,
                  const uint32_t *__restrict__ d_n, uint32_t cap, int shift, uint32_t mask,
                  const uint32_t *__restrict__ block_hist, uint32_t stride,
                  const uint32_t *__restrict__ digit_totals) {
  extern __shared__ __align__(16) unsigned char rs_smem_raw[];
  RsSmem &sm = *reinterpret_cast<RsSmem *>(rs_smem_raw);
  const uint32_t n = num_elems(d_n, cap);
  const uint32_t ntiles = (uint32_t)(((uint64_t)n + RS_TILE - 1) / RS_TILE);
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t lt = lanemask_lt();
  // global base of digit `tid`: exclusive scan of the digit totals (same for every tile)
  const uint32_t dt = tid <= mask ? digit_totals[tid] : 0;
  const uint32_t dbase = cta_excl_scan256(dt, sm.wt, lane, warp);

  for (uint32_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const uint32_t tile_n = min((uint32_t)RS_TILE, n - t * RS_TILE);
    for (int k = tid; k < RS_WARPS * 256; k += RS_THREADS) (&sm.warp_hist[0][0])[k] = 0;
    __syncthreads();

    uint32_t key[RS_ITEMS];
    uint32_t rank[RS_ITEMS];
    const uint32_t seg = t * RS_TILE + warp * (32 * RS_ITEMS);
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
      const uint32_t idx = seg + k * 32 + lane;
      key[k] = idx < n ? keys_in[idx] : 0xffffffffu;
    }
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
      const bool valid = seg + k * 32 + lane < n;
      const uint32_t d = valid ? (key[k] >> shift) & mask : 256u;
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
    for (int w = 0; w < RS_WARPS; ++w) {
      const uint32_t c = sm.warp_hist[w][tid];
      sm.warp_hist[w][tid] = cnt;
      cnt += c;
    }
    const uint32_t tstart = cta_excl_scan256(cnt, sm.wt, lane, warp);
    sm.tile_start[tid] = tstart;
    sm.out_base[tid] = tid <= mask ? dbase + block_hist[(size_t)tid * stride + t] - tstart : 0;
    __syncthreads();

    // reorder through shared memory so the global writes are runs of equal digits
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
      const uint32_t idx = seg + k * 32 + lane;
      if (idx < n) {
        const uint32_t d = (key[k] >> shift) & mask;
        const uint32_t pos = sm.tile_start[d] + sm.warp_hist[warp][d] + rank[k];
        sm.keys[pos] = key[k];
        sm.vals[pos] = vals_in ? vals_in[idx] : idx;   // first digit of sort A: value = index
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
      const uint32_t pos = k * RS_THREADS + tid;
      if (pos < tile_n) {
        const uint32_t kk = sm.keys[pos];
        const uint32_t dst = sm.out_base[(kk >> shift) & mask] + pos;
        keys_out[dst] = kk;
        vals_out[dst] = sm.vals[pos];
      }
    }
    __syncthreads();
  }
}

void sort_plan(int nbits, int &passes, int &bits_per_pass) {
  if (nbits < 1) nbits = 1;
  passes = (nbits + 7) / 8;
  bits_per_pass = (nbits + passes - 1) / passes;
}

int launch_radix_sort(SortBuffers &sb, const uint32_t *d_n, uint32_t n_upper, int begin_bit, int end_bit,
                      bool identity_vals, cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(rs_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(RsSmem));
    attr_set = true;
  }
  if (n_upper > sb.capacity) n_upper = sb.capacity;
  const uint32_t ntiles = (uint32_t)(((uint64_t)n_upper + RS_TILE - 1) / RS_TILE);
  int cur = 0;
  if (ntiles == 0) return cur;
  int passes, bpp;
  sort_plan(end_bit - begin_bit, passes, bpp);
  const uint32_t grid = ntiles < (uint32_t)RS_GRID ? ntiles : (uint32_t)RS_GRID;
  for (int p = 0; p < passes; ++p) {
    const int shift = begin_bit + p * bpp;
    const int bits = (end_bit - shift) < bpp ? (end_bit - shift) : bpp;
    const uint32_t mask = (1u << bits) - 1u;
    g_launches += 3;
    rs_hist_kernel<<<grid, RS_THREADS, 0, st>>>(sb.keys[cur], d_n, sb.capacity, shift, mask, sb.block_hist,
                                               sb.max_tiles);
    rs_scan_kernel<<<mask + 1, 256, 0, st>>>(sb.block_hist, sb.max_tiles, d_n, sb.capacity, sb.digit_totals);
    rs_scatter_kernel<<<grid, RS_THREADS, sizeof(RsSmem), st>>>(
        sb.keys[cur], (p == 0 && identity_vals) ? nullptr : sb.vals[cur], sb.keys[cur ^ 1], sb.vals[cur ^ 1], d_n,
        sb.capacity, shift, mask, sb.block_hist, sb.max_tiles, sb.digit_totals);
    cur ^= 1;
  }
  return cur;
}

// ---- K4: per-tile ranges over the sorted tile keys (counts = end - start) -------
__global__ void __launch_bounds__(256)
tile_ranges_kernel(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ d_n, uint32_t cap,
                   uint2 *__restrict__ ranges, uint32_t num_tiles) {
  const uint32_t n = num_elems(d_n, cap);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t t = keys[i];
    if (t >= num_tiles) continue;
    if (i == 0 || keys[i - 1] != t) ranges[t].x = i;
    if (i == n - 1 || keys[i + 1] != t) ranges[t].y = i + 1;
  }
}

void launch_tile_ranges(const uint32_t *keys, const uint32_t *d_n, uint32_t n_upper, uint32_t cap, uint2 *ranges,
                        uint32_t num_tiles, cudaStream_t st) {
  cudaMemsetAsync(ranges, 0, sizeof(uint2) * num_tiles, st);
  if (!n_upper) return;
  const int blocks = min(ceil_div(n_upper, 256), 148 * 8);
  ++g_launches;
  tile_ranges_kernel<<<blocks, 256, 0, st>>>(keys, d_n, cap, ranges, num_tiles);
}

}  // namespace gsplat
