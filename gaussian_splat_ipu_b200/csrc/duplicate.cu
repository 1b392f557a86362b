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

// look-back word: flag in the top two bits, 62-bit value below (one 64-bit store publishes both)
constexpr unsigned long long LB_AGGREGATE = 1ull << 62;   // value = this CTA's own total
constexpr unsigned long long LB_PREFIX = 2ull << 62;      // value = inclusive prefix up to this CTA
constexpr unsigned long long LB_VALUE = (1ull << 62) - 1;

__device__ __forceinline__ unsigned long long ld_relaxed(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed(unsigned long long *p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// One CTA = one link of the chain = K2_CHUNK consecutive positions of the depth order
// (K2_ITEMS per thread, striped so the loads coalesce and the rectangle gathers overlap).
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

// writes the `cnt` (tile, gaussian) pairs q = first, first + step, ... of one rectangle
__device__ __forceinline__ void emit_rect(uint32_t *__restrict__ keys, uint32_t *__restrict__ vals, uint32_t at,
                                          uint2 rc, uint32_t g, int tiles_across, uint32_t first, uint32_t step,
                                          uint32_t total) {
  const uint32_t c0 = rc.x & 0xffffu, w = (rc.x >> 16) - c0 + 1, r0 = rc.y & 0xffffu;
  for (uint32_t q = first; q < total; q += step) {
    const uint32_t ty = r0 + q / w, tx = c0 + q % w;
    keys[at + q] = ty * (uint32_t)tiles_across + tx;
    vals[at + q] = g;
  }
}

__global__ void __launch_bounds__(K2_THREADS)
duplicate_kernel(Projected p, const uint32_t *__restrict__ sorted_keys, const uint32_t *__restrict__ sorted_gauss,
                 uint32_t n, int tiles_across, uint32_t *__restrict__ keys, uint32_t *__restrict__ vals,
                 uint32_t capacity, unsigned long long *__restrict__ lookback, uint32_t *__restrict__ ticket,
                 uint32_t *__restrict__ total_out, uint32_t *__restrict__ overflow) {
  extern __shared__ __align__(16) unsigned char k2_smem_raw[];
  K2Smem &sm = *reinterpret_cast<K2Smem *>(k2_smem_raw);
  const uint32_t tid = threadIdx.x;
  if (tid == 0) { sm.chain = atomicAdd(ticket, 1u); sm.n_big = 0; }
  __syncthreads();
  const uint32_t b = sm.chain;
  uint32_t g[K2_ITEMS];
  uint2 rc[K2_ITEMS];
  uint32_t cnt[K2_ITEMS];
  bool live[K2_ITEMS];
#pragma unroll
  for (int i = 0; i < K2_ITEMS; ++i) {
    const uint32_t j = b * K2_CHUNK + i * K2_THREADS + tid;   // position in depth order
    live[i] = j < n && sorted_keys[j] != DEPTH_KEY_CULLED;
    g[i] = live[i] ? sorted_gauss[j] : 0u;
  }
#pragma unroll
  for (int i = 0; i < K2_ITEMS; ++i) {
    rc[i] = live[i] ? __ldg(&p.rect[g[i]]) : make_uint2(0u, 0u);
    const uint32_t w = (rc[i].x >> 16) - (rc[i].x & 0xffffu) + 1, h = (rc[i].y >> 16) - (rc[i].y & 0xffffu) + 1;
    cnt[i] = live[i] ? w * h : 0u;
    sm.prefix[i * K2_THREADS + tid] = cnt[i];
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
    unsigned long long excl = 0;
    if (b == 0) {
      if (tid == 0) st_relaxed(&lookback[0], LB_PREFIX | total);
    } else {
      if (tid == 0) st_relaxed(&lookback[b], LB_AGGREGATE | total);
      int pos = (int)b - 1;
      while (true) {   // one warp inspects 32 predecessors at a time
        const int q = pos - (int)tid;
        unsigned long long w = LB_PREFIX;   // positions before the chain start count as prefix 0
        if (q >= 0) {
          do { w = ld_relaxed(&lookback[q]); } while ((w >> 62) == 0);
        }
        const uint32_t is_prefix = __ballot_sync(0xffffffffu, (w >> 62) == 2);
        const int first = is_prefix ? __ffs(is_prefix) - 1 : 32;   // nearest predecessor with a full prefix
        unsigned long long v = (int)tid <= first ? (w & LB_VALUE) : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        excl += v;
        if (is_prefix) break;
        pos -= 32;
      }
      if (tid == 0) st_relaxed(&lookback[b], LB_PREFIX | (excl + total));
    }
    if (tid == 0) {
      const unsigned long long incl = excl + total;
      if (b == gridDim.x - 1) *total_out = incl > 0xffffffffull ? 0xffffffffu : (uint32_t)incl;
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
  // large ones are queued and written by the whole CTA so they do not serialise one thread.
#pragma unroll
  for (int i = 0; i < K2_ITEMS; ++i) {
    if (cnt[i] == 0) continue;
    const uint32_t e = i * K2_THREADS + tid;
    if (cnt[i] <= K2_SMALL) {   // row-major walk of the rectangle without divisions
      const uint32_t c0 = rc[i].x & 0xffffu, c1 = rc[i].x >> 16;
      uint32_t at = base + sm.prefix[e], tx = c0, row = (rc[i].y & 0xffffu) * (uint32_t)tiles_across;
      for (uint32_t q = 0; q < cnt[i]; ++q) {
        keys[at] = row + tx;
        vals[at] = g[i];
        ++at;
        if (tx == c1) { tx = c0; row += (uint32_t)tiles_across; } else { ++tx; }
      }
    } else {
      const uint32_t slot = atomicAdd(&sm.n_big, 1u);
      sm.big[slot] = e; sm.big_rect[slot] = rc[i]; sm.big_gauss[slot] = g[i];
    }
  }
  __syncthreads();
  const uint32_t n_big = sm.n_big;
  for (uint32_t s = 0; s < n_big; ++s) {
    const uint2 r = sm.big_rect[s];
    const uint32_t w = (r.x >> 16) - (r.x & 0xffffu) + 1, h = (r.y >> 16) - (r.y & 0xffffu) + 1;
    emit_rect(keys, vals, base + sm.prefix[sm.big[s]], r, sm.big_gauss[s], tiles_across, tid, K2_THREADS, w * h);
  }
}

void launch_duplicate(const Projected &p, const uint32_t *sorted_keys, const uint32_t *sorted_gauss, uint32_t n,
                      int tiles_across, uint32_t *keys, uint32_t *vals, uint32_t capacity,
                      unsigned long long *lookback, uint32_t *ticket, uint32_t *total, uint32_t *overflow,
                      cudaStream_t st) {
  if (!n) return;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(duplicate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(K2Smem));
    attr_set = true;
  }
  ++g_launches;
  duplicate_kernel<<<ceil_div(n, K2_CHUNK), K2_THREADS, sizeof(K2Smem), st>>>(p, sorted_keys, sorted_gauss, n, tiles_across, keys,
                                                                  vals, capacity, lookback, ticket, total, overflow);
}

}  // namespace gsplat
