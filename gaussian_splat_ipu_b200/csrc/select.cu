// select.cu — depth-slice selection: which Gaussians enter sort A / K2 in each depth slice.
//
// K5 stops a pixel at the first Gaussian that would push its transmittance under 1e-4
// (codelets/splat/codelets.cpp:405-408), so once every pixel of a tile has stopped nothing further
// down the tile's z-ordered list (sortBuffer, codelets.cpp:303-356) can change the image.  The
// frame is therefore processed in a few slices of increasing depth.  Slice s takes the Gaussians
// whose depth key lies in [tau[s-1], tau[s]) — and, from the second slice on, whose rectangle
// still covers a tile that has not stopped — sorts only those by depth, bins and blends them, and
// K5 raises tile_done for tiles that stopped.  Every tile's list is an exact prefix of the
// reference's list (complete unless the tile stopped), so the framebuffer is bit-identical to the
// one-slice run, while the bulk of a dense scene — Gaussians buried behind saturated tiles — is
// never sorted or duplicated at all.
//
// The thresholds tau are depth-key quantiles of a strided sample of this frame's keys: ANY
// thresholds give the same image (they only decide where the slices are cut), so an approximate
// quantile is enough and no state is carried from frame to frame.
#include "gsplat_internal.h"

namespace gsplat {

// ---- thresholds: quantiles of a 4096-key sample, one CTA ----------------------------------------
// fractions[s] -> tau[s] = the sample key of rank fraction * (drawn sample keys), found by a
// four-round radix select over the register-resident sample.
constexpr int QS_THREADS = 1024;
constexpr int QS_PER_THREAD = 4;
constexpr int QS_SAMPLE = QS_THREADS * QS_PER_THREAD;

__global__ void __launch_bounds__(QS_THREADS)
depth_quantile_kernel(const uint32_t *__restrict__ depth_key, uint32_t n, SliceFractions fr, uint32_t *__restrict__ tau) {
  // radix select, 8 bits per round: histogram of the sample keys that share the prefix found so
  // far, then the bin in which the running count crosses the target rank
  __shared__ uint32_t hist[256];
  __shared__ uint32_t sh_prefix, sh_rank, sh_drawn;
  const uint32_t tid = threadIdx.x;
  uint32_t key[QS_PER_THREAD];
#pragma unroll
  for (int k = 0; k < QS_PER_THREAD; ++k) {
    const uint32_t j = tid * QS_PER_THREAD + k;
    const uint64_t i = n > QS_SAMPLE ? (uint64_t)j * n / QS_SAMPLE : j;
    key[k] = i < n ? depth_key[i] : DEPTH_KEY_CULLED;
  }
  {  // drawn sample keys
    uint32_t c = 0;
#pragma unroll
    for (int k = 0; k < QS_PER_THREAD; ++k) c += key[k] != DEPTH_KEY_CULLED ? 1u : 0u;
    if (tid == 0) sh_drawn = 0;
    __syncthreads();
    c = __reduce_add_sync(0xffffffffu, c);
    if ((tid & 31) == 0 && c) atomicAdd(&sh_drawn, c);
    __syncthreads();
  }
  const uint32_t drawn = sh_drawn;
  for (int s = 0; s < fr.count; ++s) {
    // tau = the sample key of rank `target` among the drawn ones (0-based); keys below it form the slice
    if (tid == 0) { sh_prefix = 0; sh_rank = min((uint32_t)((double)drawn * fr.f[s]), drawn ? drawn - 1 : 0u); }
    for (int shift = 24; shift >= 0; shift -= 8) {
      if (tid < 256) hist[tid] = 0;
      __syncthreads();
      const uint32_t prefix = sh_prefix, hi_mask = shift == 24 ? 0u : 0xffffffffu << (shift + 8);
#pragma unroll
      for (int k = 0; k < QS_PER_THREAD; ++k)
        if (key[k] != DEPTH_KEY_CULLED && (key[k] & hi_mask) == prefix) atomicAdd(&hist[(key[k] >> shift) & 255u], 1u);
      __syncthreads();
      if (tid < 32) {   // one warp: find the bin where the running count passes the rank
        uint32_t v[8], sum = 0;
#pragma unroll
        for (int q = 0; q < 8; ++q) { v[q] = hist[tid * 8 + q]; sum += v[q]; }
        uint32_t inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
          if (tid >= o) inc += t;
        }
        const uint32_t rank = sh_rank;
        uint32_t run = inc - sum;
        if (rank >= run && rank < inc) {   // exactly one lane
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            if (rank >= run && rank < run + v[q]) { sh_prefix = prefix | ((tid * 8 + q) << shift); sh_rank = rank - run; }
            run += v[q];
          }
        }
      }
      __syncthreads();
    }
    if (tid == 0) tau[s] = drawn ? sh_prefix : DEPTH_KEY_CULLED;
    __syncthreads();
  }
}

void launch_depth_quantiles(const uint32_t *depth_key, uint32_t n, const SliceFractions &fr, uint32_t *tau,
                            cudaStream_t st) {
  if (!fr.count) return;
  ++g_launches;
  depth_quantile_kernel<<<1, QS_THREADS, 0, st>>>(depth_key, n, fr, tau);
}

// ---- order-preserving selection (stream compaction with a chained scan) ---------------------------
// One CTA = SEL_CHUNK consecutive Gaussians; positions come from warp ballots plus a scan of the
// per-(item row, warp) counts, so the selected (key, index) pairs come out in index order: the stable sort that follows then
// breaks depth ties by gid exactly as the one-slice path does.
constexpr int SEL_THREADS = 256;
constexpr int SEL_ITEMS = 16;
constexpr int SEL_CHUNK = SEL_THREADS * SEL_ITEMS;

template <bool CHECK_DONE>
__global__ void __launch_bounds__(SEL_THREADS)
select_kernel(const uint32_t *__restrict__ depth_key, const uint2 *__restrict__ rect, uint32_t n,
              const uint32_t *__restrict__ tau_lo /* nullptr: 0 */, const uint32_t *__restrict__ tau_hi /* nullptr: all drawn */,
              int tiles_across, const uint32_t *__restrict__ live_sat, uint32_t *__restrict__ keys_out,
              uint32_t *__restrict__ idx_out, unsigned long long *__restrict__ lookback, uint32_t *__restrict__ ticket,
              uint32_t *__restrict__ total_out, uint32_t *__restrict__ key_range /* {max key, max ~key} */) {
  __shared__ uint32_t wc[SEL_ITEMS * (SEL_THREADS / 32) + 1];
  __shared__ uint32_t wmax[2 * (SEL_THREADS / 32)];
  __shared__ uint32_t chain_s;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) chain_s = atomicAdd(ticket, 1u);
  __syncthreads();
  const uint32_t b = chain_s;
  const uint32_t lo = tau_lo ? *tau_lo : 0u, hi = tau_hi ? *tau_hi : DEPTH_KEY_CULLED;
  // striped: item k of thread tid is Gaussian base + k * SEL_THREADS + tid, so every load of a warp
  // is one contiguous run; output order = (k, tid) = index order
  const uint32_t base_i = b * SEL_CHUNK;
  uint32_t key[SEL_ITEMS];
#pragma unroll
  for (int k = 0; k < SEL_ITEMS; ++k) {
    const uint32_t i = base_i + k * SEL_THREADS + tid;
    key[k] = i < n ? depth_key[i] : DEPTH_KEY_CULLED;
  }
  uint32_t sel = 0;   // bit k: item k is selected
  uint32_t kmax = 0u, kinv = 0u;
#pragma unroll
  for (int k = 0; k < SEL_ITEMS; ++k) {
    bool take = key[k] >= lo && key[k] < hi;   // hi <= DEPTH_KEY_CULLED: culled Gaussians never pass
    if (CHECK_DONE && take) {   // keep it only if its rectangle still covers a tile that is being blended
      const uint2 rc = rect[base_i + k * SEL_THREADS + tid];
      take = live_tiles_in_rect(live_sat, tiles_across, rc.x & 0xffffu, rc.x >> 16, rc.y & 0xffffu, rc.y >> 16) != 0;
    }
    sel |= take ? (1u << k) : 0u;
    const uint32_t m = __ballot_sync(0xffffffffu, take);
    if (lane == 0) wc[k * (SEL_THREADS / 32) + warp] = __popc(m);
    if (take) { kmax = max(kmax, key[k]); kinv = max(kinv, ~key[k]); }
  }
  // range of the selected keys: the depth sort skips the digits above the bits in which they differ
  kmax = __reduce_max_sync(0xffffffffu, kmax);
  kinv = __reduce_max_sync(0xffffffffu, kinv);
  if (lane == 0) { wmax[warp] = kmax; wmax[SEL_THREADS / 32 + warp] = kinv; }
  __syncthreads();
  if (tid < 2 * (SEL_THREADS / 32)) {   // lanes 0-7: max key, lanes 8-15: max ~key
    uint32_t v = wmax[tid];
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0x0000ffffu, v, o));
    if ((tid & 7u) == 0 && v) atomicMax(&key_range[tid >> 3], v);
  }
  // exclusive scan of the SEL_ITEMS x warps counts in (k, warp) order: one warp, 4 entries per lane
  constexpr int NWC = SEL_ITEMS * (SEL_THREADS / 32);
  static_assert(NWC == 128, "one warp scans 4 entries per lane");
  if (warp == 0) {
    uint32_t v[4], sum = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) { v[q] = wc[lane * 4 + q]; sum += v[q]; }
    uint32_t inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    uint32_t run = inc - sum;
#pragma unroll
    for (int q = 0; q < 4; ++q) { wc[lane * 4 + q] = run; run += v[q]; }
    if (lane == 31) wc[NWC] = inc;
  }
  __syncthreads();
  const uint32_t total = wc[NWC];
  // chained scan over CTAs (same scheme as K2's)
  if (warp == 0) {
    const unsigned long long excl = lb_exclusive_prefix(lookback, b, total, lane);
    if (lane == 0) {
      chain_s = (uint32_t)excl;
      if (b == gridDim.x - 1) *total_out = (uint32_t)(excl + total);
    }
  }
  __syncthreads();
  const uint32_t cbase = chain_s;
#pragma unroll
  for (int k = 0; k < SEL_ITEMS; ++k) {
    const bool take = sel >> k & 1u;
    const uint32_t m = __ballot_sync(0xffffffffu, take);
    if (take) {
      const uint32_t at = cbase + wc[k * (SEL_THREADS / 32) + warp] + __popc(m & ((1u << lane) - 1u));
      keys_out[at] = key[k];
      idx_out[at] = base_i + k * SEL_THREADS + tid;
    }
  }
}

void launch_select(const Projected &p, uint32_t n, const uint32_t *tau_lo, const uint32_t *tau_hi, int tiles_across,
                   const uint32_t *live_sat, uint32_t *keys_out, uint32_t *idx_out, unsigned long long *lookback,
                   uint32_t *ticket, uint32_t *total_out, uint32_t *key_range, cudaStream_t st) {
  if (!n) return;
  ++g_launches;
  const int grid = ceil_div(n, SEL_CHUNK);
  if (live_sat)
    select_kernel<true><<<grid, SEL_THREADS, 0, st>>>(p.depth_key, p.rect, n, tau_lo, tau_hi, tiles_across, live_sat,
                                                      keys_out, idx_out, lookback, ticket, total_out, key_range);
  else
    select_kernel<false><<<grid, SEL_THREADS, 0, st>>>(p.depth_key, p.rect, n, tau_lo, tau_hi, tiles_across, nullptr,
                                                       keys_out, idx_out, lookback, ticket, total_out, key_range);
}

}  // namespace gsplat
