// radix_sort.cu — K3: hand-written device-wide stable LSD radix sort of (u32 key, u32 value)
// pairs, and K4 (per-tile ranges).
//
// The reference orders each tile's list with a per-tile quicksort by clip z
// (codelets/splat/codelets.cpp:303-356) after its lattice has routed Gaussians to tiles.  Here the
// same order — (tile, z ascending, gid ascending) — comes from TWO uses of this one sort:
//   sort A  the N Gaussians by their 32-bit depth key (5 digits of 7/6 bits), values = Gaussian index
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
// per element per digit; digits are at most 7 bits wide (8 for the small per-slice depth sorts,
// whose 128 KB of counters leave one CTA per SM).
#include "gsplat_internal.h"
#include <stdlib.h>

namespace gsplat {

__device__ __forceinline__ uint32_t lanemask_lt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

__device__ __forceinline__ uint32_t num_elems(const uint32_t *d_n, uint32_t cap) {
  // a count beyond the capacity means the producer overflowed the buffer and wrote an incomplete
  // array (the host re-runs the frame): sort nothing rather than garbage
  const uint32_t n = *d_n;
  return n <= cap ? n : 0u;
}

// Elements a pass has to move: none when its digit lies above the bits in which the keys differ
// (launch_radix_sort's key_range): the pass is skipped on the device and the data stays where it is.
__device__ __forceinline__ uint32_t pass_elems(const uint32_t *d_n, uint32_t cap, const uint32_t *key_range, int shift) {
  if (key_range && shift >= (int)key_range_bits(key_range)) return 0u;
  return num_elems(d_n, cap);
}

constexpr int RS_WARPS = RS_THREADS / 32;

// Ranking uses per-thread digit counters in shared memory (no warp votes, no atomics): every
// thread owns RS_ITEMS CONSECUTIVE elements of the tile and bumps a private 16-bit counter per
// digit; a raking scan over all counters in (digit, thread) order then turns "count so far in my
// thread" into the element's stable position inside the tile.  Two digits share a 32-bit word
// (digit d and d + LANES, LANES = radix / 2), so the scan adds whole words: no half can overflow
// because a tile has 4096 < 65536 elements.  Word (lane, thread) lives at PAD(lane * 256 + thread);
// the padding (two words per 32) keeps the per-thread bumps and the 64-bit raking loads
// bank-conflict free and 8-byte aligned.
__host__ __device__ constexpr uint32_t rs_pad(uint32_t i) { return i + ((i >> 5) << 1); }
template <int BITS> struct RsGeom {
  static constexpr int LANES_LOG = BITS > 1 ? BITS - 1 : 0;
  static constexpr uint32_t LANES = 1u << LANES_LOG;          // counter words per thread
  static constexpr uint32_t MASK = (1u << BITS) - 1u;
  static constexpr uint32_t WORDS = rs_pad(LANES * RS_THREADS);
  static constexpr uint32_t LANE_STRIDE = rs_pad(RS_THREADS) * 4u;   // bytes between counter lanes
};
__host__ __device__ constexpr uint32_t rs_counter_words(int bits) { return rs_pad((uint32_t)RS_THREADS << (bits > 1 ? bits - 1 : 0)); }

// exclusive scan of one value per thread over the 256-thread CTA; *total = sum (all threads)
__device__ __forceinline__ uint32_t cta_excl_scan256(uint32_t v, uint32_t *wt, uint32_t lane, uint32_t warp,
                                                     uint32_t *total = nullptr) {
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
    if (lane == RS_WARPS - 1) wt[RS_WARPS] = ti;
  }
  __syncthreads();
  const uint32_t r = wt[warp] + inc - v;
  if (total) *total = wt[RS_WARPS];
  __syncthreads();
  return r;
}

// Loads the RS_ITEMS consecutive keys this thread owns (blocked arrangement); elements past n get
// the all-ones key, which ranks after every valid element of the tile.
__device__ __forceinline__ void rs_load_blocked(const uint32_t *__restrict__ src, uint32_t first, uint32_t n,
                                                uint32_t fill, uint32_t (&out)[RS_ITEMS]) {
  if (first + RS_ITEMS <= n) {
    const uint4 *p = reinterpret_cast<const uint4 *>(src + first);
#pragma unroll
    for (int q = 0; q < RS_ITEMS / 4; ++q) {
      const uint4 v = p[q];
      out[q * 4 + 0] = v.x; out[q * 4 + 1] = v.y; out[q * 4 + 2] = v.z; out[q * 4 + 3] = v.w;
    }
  } else {
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) out[k] = first + k < n ? src[first + k] : fill;
  }
}

// byte offset of the 16-bit counter of digit d for the thread whose padded byte offset is tid_off
template <int BITS>
__device__ __forceinline__ uint32_t rs_counter_off(uint32_t d, uint32_t tid_off) {
  using G = RsGeom<BITS>;
  return (d & (G::LANES - 1u)) * G::LANE_STRIDE + (d >> G::LANES_LOG) * 2u + tid_off;
}

// Sum (SCAN = false) or in-place exclusive scan seeded with `run` (SCAN = true) of the COUNT
// consecutive counter words starting at word index w0 (w0 a multiple of COUNT or of 32).
template <uint32_t COUNT, bool SCAN>
__device__ __forceinline__ uint32_t rs_rake(uint32_t *cnt, uint32_t w0, uint32_t run) {
  if constexpr (COUNT == 1) {
    const uint32_t a = rs_pad(w0), w = cnt[a];
    if (SCAN) cnt[a] = run;
    return run + w;
  } else {
#pragma unroll
    for (uint32_t seg = 0; seg < COUNT; seg += 32) {
      uint2 *p = reinterpret_cast<uint2 *>(cnt + rs_pad(w0 + seg));
      constexpr uint32_t PAIRS = (COUNT < 32 ? COUNT : 32) / 2;
#pragma unroll
      for (uint32_t q = 0; q < PAIRS; ++q) {
        const uint2 w = p[q];
        if (SCAN) p[q] = make_uint2(run, run + w.x);
        run += w.x + w.y;
      }
    }
    return run;
  }
}

// ---- digit histogram of every tile -----------------------------------------------------------
template <int BITS>
__global__ void __launch_bounds__(RS_THREADS)
rs_hist_kernel(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ d_n, uint32_t cap, int shift,
               uint32_t *__restrict__ block_hist, uint32_t stride, const uint32_t *__restrict__ key_range) {
  using G = RsGeom<BITS>;
  extern __shared__ __align__(16) uint32_t rs_cnt[];
  const uint32_t n = pass_elems(d_n, cap, key_range, shift);
  const uint32_t ntiles = (uint32_t)(((uint64_t)n + RS_TILE - 1) / RS_TILE);
  const uint32_t tid = threadIdx.x;
  const uint32_t tid_off = rs_pad(tid) * 4u;
  // reduction mapping: thread = (counter lane, part); each sums SPAN threads' packed words
  constexpr uint32_t PARTS = RS_THREADS / G::LANES, SPAN = G::LANES;   // PARTS * SPAN == RS_THREADS
  const uint32_t r_lane = tid / PARTS, r_part = tid % PARTS;
  unsigned char *cbytes = reinterpret_cast<unsigned char *>(rs_cnt);
  for (uint32_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
    for (uint32_t k = tid * 2; k < G::WORDS; k += RS_THREADS * 2) *reinterpret_cast<uint2 *>(rs_cnt + k) = make_uint2(0, 0);
    __syncthreads();
    uint32_t key[RS_ITEMS];
    const uint32_t first = t * RS_TILE + tid * RS_ITEMS;
    rs_load_blocked(keys, first, n, 0xffffffffu, key);
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
      if (first + k < n) {
        unsigned short *c = reinterpret_cast<unsigned short *>(
            cbytes + rs_counter_off<BITS>((key[k] >> shift) & G::MASK, tid_off));
        *c = *c + 1;
      }
    }
    __syncthreads();
    // packed sum: low half = digit r_lane, high half = digit r_lane + LANES
    uint32_t sum = rs_rake<SPAN, false>(rs_cnt, r_lane * RS_THREADS + r_part * SPAN, 0u);
#pragma unroll
    for (uint32_t o = 1; o < PARTS && o < 32; o <<= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if constexpr (PARTS > 32) {   // narrow digits: finish across warps through shared memory
      __syncthreads();
      if ((tid & 31) == 0) rs_cnt[tid >> 5] = sum;
      __syncthreads();
      sum = 0;
      if (r_part == 0) for (uint32_t w = 0; w < PARTS / 32; ++w) sum += rs_cnt[r_lane * (PARTS / 32) + w];
    }
    if (r_part == 0) {
      block_hist[(size_t)r_lane * stride + t] = sum & 0xffffu;
      block_hist[(size_t)(r_lane + G::LANES) * stride + t] = sum >> 16;
    }
    __syncthreads();
  }
}

// ---- exclusive scan of each digit's row over tiles; digit totals -------------------------------
__global__ void __launch_bounds__(256)
rs_scan_kernel(uint32_t *__restrict__ block_hist, uint32_t stride, const uint32_t *__restrict__ d_n, uint32_t cap,
               uint32_t *__restrict__ digit_totals, const uint32_t *__restrict__ key_range, int shift) {
  __shared__ uint32_t wt[256 / 32 + 1];
  const uint32_t n = pass_elems(d_n, cap, key_range, shift);
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
// Dynamic shared memory: the counter words, re-used as the (key, value) staging area once every
// element knows its position; then out_base[256] and the scan scratch.
struct RsTail {
  uint32_t out_base[256];
  uint32_t wt[RS_WARPS + 1];
  uint32_t tstart[257];     // rs_sweep_kernel: per-digit start inside the tile (+ the tile's length)
  uint32_t ticket;          // rs_sweep_kernel: the tile this CTA works on
};
__host__ __device__ constexpr size_t rs_scatter_smem(int bits) {
  const size_t counters = (size_t)rs_counter_words(bits) * 4, staging = (size_t)RS_TILE * 8;
  return (counters > staging ? counters : staging) + sizeof(RsTail);
}

template <int BITS>
__global__ void __launch_bounds__(RS_THREADS, 3)
rs_scatter_kernel(const uint32_t *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
                  uint32_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out,
                  const uint32_t *__restrict__ d_n, uint32_t cap, int shift,
                  const uint32_t *__restrict__ block_hist, uint32_t stride,
                  const uint32_t *__restrict__ digit_totals, const uint32_t *__restrict__ key_range) {
  using G = RsGeom<BITS>;
  extern __shared__ __align__(16) uint32_t rs_cnt[];
  unsigned char *cbytes = reinterpret_cast<unsigned char *>(rs_cnt);
  uint2 *stage = reinterpret_cast<uint2 *>(rs_cnt);
  RsTail &tail = *reinterpret_cast<RsTail *>(cbytes + (rs_scatter_smem(BITS) - sizeof(RsTail)));
  const uint32_t n = pass_elems(d_n, cap, key_range, shift);
  const uint32_t ntiles = (uint32_t)(((uint64_t)n + RS_TILE - 1) / RS_TILE);
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t tid_off = rs_pad(tid) * 4u;
  // global base of digit `tid`: exclusive scan of the digit totals (same for every tile)
  const uint32_t dt = tid <= G::MASK ? digit_totals[tid] : 0;
  const uint32_t dbase = cta_excl_scan256(dt, tail.wt, lane, warp);

  for (uint32_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const uint32_t tile_n = min((uint32_t)RS_TILE, n - t * RS_TILE);
    for (uint32_t k = tid * 2; k < G::WORDS; k += RS_THREADS * 2) *reinterpret_cast<uint2 *>(rs_cnt + k) = make_uint2(0, 0);
    // this tile's per-digit offset among all tiles (independent of the ranking: fetch it early)
    const uint32_t tile_digit_off = tid <= G::MASK ? block_hist[(size_t)tid * stride + t] : 0;
    uint32_t key[RS_ITEMS], val[RS_ITEMS], pos[RS_ITEMS];
    const uint32_t first = t * RS_TILE + tid * RS_ITEMS;
    rs_load_blocked(keys_in, first, n, 0xffffffffu, key);
    if (vals_in) rs_load_blocked(vals_in, first, n, 0u, val);
    else {   // first digit of sort A: value = element index
#pragma unroll
      for (int k = 0; k < RS_ITEMS; ++k) val[k] = first + k;
    }
    __syncthreads();
    // count: pos[k] = how many earlier elements of THIS thread have the same digit
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
      unsigned short *c = reinterpret_cast<unsigned short *>(
          cbytes + rs_counter_off<BITS>((key[k] >> shift) & G::MASK, tid_off));
      const uint32_t r = *c;
      *c = (unsigned short)(r + 1);
      pos[k] = r;
    }
    __syncthreads();

    // raking scan over all counter words in (counter lane, thread) order: thread `tid` owns the
    // LANES consecutive words starting at tid * LANES
    {
      const uint32_t sum = rs_rake<G::LANES, false>(rs_cnt, tid * G::LANES, 0u);
      uint32_t total;
      uint32_t run = cta_excl_scan256(sum, tail.wt, lane, warp, &total);
      run += total << 16;   // every low-half digit (d < LANES) precedes every high-half digit
      rs_rake<G::LANES, true>(rs_cnt, tid * G::LANES, run);
    }
    __syncthreads();
    // position inside the digit-sorted tile; per-digit start of the tile -> global destination base
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
      const unsigned short *c = reinterpret_cast<const unsigned short *>(
          cbytes + rs_counter_off<BITS>((key[k] >> shift) & G::MASK, tid_off));
      pos[k] += *c;
    }
    if (tid <= G::MASK) {
      const uint32_t tstart = *reinterpret_cast<const unsigned short *>(cbytes + rs_counter_off<BITS>(tid, 0u));
      tail.out_base[tid] = dbase + tile_digit_off - tstart;
    }
    __syncthreads();   // counters are dead: the same memory now stages the sorted tile
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) stage[pos[k]] = make_uint2(key[k], val[k]);
    __syncthreads();
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
      const uint32_t q = k * RS_THREADS + tid;
      if (q < tile_n) {
        const uint2 kv = stage[q];
        const uint32_t dst = tail.out_base[(kv.x >> shift) & G::MASK] + q;
        keys_out[dst] = kv.x;
        vals_out[dst] = kv.y;
      }
    }
    __syncthreads();
  }
}

// ---- single-sweep passes ---------------------------------------------------------------------------
// The three-kernel pass above costs three launches and three dependent round trips per digit; the
// frame's two sorts per depth slice are 42 of its 57 launches, each a few microseconds long.  The
// single-sweep form does a digit in ONE launch:
//   rs_hist_all_kernel   once per sort: the digit totals of EVERY pass (the keys do not change
//                        between passes, only their order), one read of the keys;
//   rs_sweep_kernel      per digit: ranks a tile exactly as rs_scatter_kernel does, publishes the tile's
//                        per-digit counts and obtains its per-digit offsets from the tiles before it by
//                        decoupled look-back (one 64-bit word per (tile, digit): sort serial | pass,
//                        flag, value — a word of another sort or pass reads as "not there yet", so
//                        nothing is cleared between sorts).  Tiles are handed out by a ticket, so every
//                        tile a CTA waits for belongs to a CTA that is already running.
// Stable (tiles are ordered by the look-back, elements inside a tile by the ranking), so the result
// is bit for bit the three-kernel sort's.  The control block cleans itself: the last CTA of a pass
// to leave zeroes that pass's totals, ticket and exit counter.
constexpr uint32_t SWEEP_FLAG_AGG = 1u, SWEEP_FLAG_INC = 2u;   // tile's own count / inclusive prefix
constexpr int SWEEP_LOOKBACK = 8;                               // tiles examined per look-back step (loads in flight)
__device__ __forceinline__ unsigned long long sweep_word(uint32_t tag, uint32_t flag, uint32_t value) {
  return ((unsigned long long)tag << 34) | ((unsigned long long)flag << 32) | value;
}
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long *p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Digit totals of all passes.  Shared-memory atomics; a thread merges runs of equal digits among its
// RS_ITEMS consecutive keys first (sort B's keys — tile ids emitted Gaussian by Gaussian — and the top
// digit of the depth keys are long runs of one value).  Passes above the key range are skipped, as
// rs_sweep_kernel skips them.
__global__ void __launch_bounds__(RS_THREADS)
rs_hist_all_kernel(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ d_n, uint32_t cap, int begin_bit,
                   int end_bit, int bpp, int passes, uint32_t *__restrict__ ctl, const uint32_t *__restrict__ key_range,
                   unsigned long long *__restrict__ tile_state, size_t tile_words, uint32_t wrap_mask) {
  __shared__ uint32_t hist[RS_SWEEP_MAX_PASSES * 256];
  __shared__ uint32_t wrapped;
  const uint32_t n = num_elems(d_n, cap);
  if (n == 0) return;
  const uint32_t tid = threadIdx.x;
  if (blockIdx.x == 0) {
    // the sort's serial number, read by the passes that follow.  The look-back words carry it in 27
    // bits (tag = serial * 8 + pass, 30 bits): long before a tag can come round again — every
    // wrap_mask + 1 = 2^26 sorts, about a day of rendering — the words are wiped, here, before any
    // pass of this sort runs, so a word of a sort 2^27 sorts ago can never be taken for a fresh one.
    if (tid == 0) {
      const uint32_t serial = ctl[RS_SWEEP_SERIAL] + 1u;
      ctl[RS_SWEEP_SERIAL] = serial;
      wrapped = (serial & wrap_mask) == 0u ? 1u : 0u;
    }
    __syncthreads();
    if (wrapped)
      for (size_t i = tid; i < tile_words; i += RS_THREADS) tile_state[i] = 0ull;
  }
  const uint32_t ntiles = (uint32_t)(((uint64_t)n + RS_TILE - 1) / RS_TILE);
  if (blockIdx.x >= ntiles) return;
  const uint32_t kbits = key_range ? key_range_bits(key_range) : 32u;
  for (uint32_t k = tid; k < (uint32_t)passes * 256u; k += RS_THREADS) hist[k] = 0;
  __syncthreads();
  for (uint32_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
    uint32_t key[RS_ITEMS];
    const uint32_t first = t * RS_TILE + tid * RS_ITEMS;
    rs_load_blocked(keys, first, n, 0xffffffffu, key);
    const uint32_t valid = first < n ? min((uint32_t)RS_ITEMS, n - first) : 0u;
    for (int p = 0; p < passes; ++p) {
      const int shift = begin_bit + p * bpp;
      if (shift >= (int)kbits) break;
      const int bits = (end_bit - shift) < bpp ? (end_bit - shift) : bpp;
      const uint32_t mask = (1u << bits) - 1u;
      uint32_t run_digit = 0, run = 0;
#pragma unroll
      for (int k = 0; k < RS_ITEMS; ++k) {
        if ((uint32_t)k < valid) {
          const uint32_t d = (key[k] >> shift) & mask;
          if (run && d != run_digit) { atomicAdd(&hist[p * 256 + run_digit], run); run = 0; }
          run_digit = d; ++run;
        }
      }
      if (run) atomicAdd(&hist[p * 256 + run_digit], run);
    }
  }
  __syncthreads();
  for (uint32_t k = tid; k < (uint32_t)passes * 256u; k += RS_THREADS) {
    const uint32_t v = hist[k];
    if (v) atomicAdd(&ctl[RS_SWEEP_HIST + k], v);
  }
}

template <int BITS>
__global__ void __launch_bounds__(RS_THREADS, 3)
rs_sweep_kernel(const uint32_t *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
                uint32_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out,
                const uint32_t *__restrict__ d_n, uint32_t cap, int shift, int pass,
                unsigned long long *__restrict__ tile_state, uint32_t *__restrict__ ctl,
                const uint32_t *__restrict__ key_range, uint32_t backoff_ns) {
  using G = RsGeom<BITS>;
  constexpr uint32_t R = 1u << BITS;
  extern __shared__ __align__(16) uint32_t rs_cnt[];
  unsigned char *cbytes = reinterpret_cast<unsigned char *>(rs_cnt);
  uint2 *stage = reinterpret_cast<uint2 *>(rs_cnt);
  RsTail &tail = *reinterpret_cast<RsTail *>(cbytes + (rs_scatter_smem(BITS) - sizeof(RsTail)));
  const uint32_t n = pass_elems(d_n, cap, key_range, shift);
  if (n == 0) return;                       // skipped pass: its control words were never touched
  const uint32_t ntiles = (uint32_t)(((uint64_t)n + RS_TILE - 1) / RS_TILE);
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t tid_off = rs_pad(tid) * 4u;
  // global base of digit `tid`: exclusive scan of this pass's digit totals (rs_hist_all_kernel)
  const uint32_t dt = tid <= G::MASK ? ctl[RS_SWEEP_HIST + pass * 256 + tid] : 0;
  const uint32_t dbase = cta_excl_scan256(dt, tail.wt, lane, warp);
  // words of this sort and pass only: anything else in tile_state reads as "not published yet"
  const uint32_t tag = (ctl[RS_SWEEP_SERIAL] * (uint32_t)RS_SWEEP_MAX_PASSES + (uint32_t)pass) & 0x3fffffffu;

  for (;;) {
    if (tid == 0) tail.ticket = atomicAdd(&ctl[RS_SWEEP_TICKET + pass], 1u);
    for (uint32_t k = tid * 2; k < G::WORDS; k += RS_THREADS * 2) *reinterpret_cast<uint2 *>(rs_cnt + k) = make_uint2(0, 0);
    __syncthreads();
    const uint32_t t = tail.ticket;
    if (t >= ntiles) break;
    const uint32_t tile_n = min((uint32_t)RS_TILE, n - t * RS_TILE);
    uint32_t key[RS_ITEMS], val[RS_ITEMS], pos[RS_ITEMS];
    const uint32_t first = t * RS_TILE + tid * RS_ITEMS;
    rs_load_blocked(keys_in, first, n, 0xffffffffu, key);
    if (vals_in) rs_load_blocked(vals_in, first, n, 0u, val);
    else {   // first digit of sort A: value = element index
#pragma unroll
      for (int k = 0; k < RS_ITEMS; ++k) val[k] = first + k;
    }
    // count: pos[k] = how many earlier elements of THIS thread have the same digit
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
      unsigned short *c = reinterpret_cast<unsigned short *>(
          cbytes + rs_counter_off<BITS>((key[k] >> shift) & G::MASK, tid_off));
      const uint32_t r = *c;
      *c = (unsigned short)(r + 1);
      pos[k] = r;
    }
    __syncthreads();
    {   // raking scan over all counter words, as in rs_scatter_kernel
      const uint32_t sum = rs_rake<G::LANES, false>(rs_cnt, tid * G::LANES, 0u);
      uint32_t total;
      uint32_t run = cta_excl_scan256(sum, tail.wt, lane, warp, &total);
      run += total << 16;
      rs_rake<G::LANES, true>(rs_cnt, tid * G::LANES, run);
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
      const unsigned short *c = reinterpret_cast<const unsigned short *>(
          cbytes + rs_counter_off<BITS>((key[k] >> shift) & G::MASK, tid_off));
      pos[k] += *c;
    }
    if (tid <= G::MASK) tail.tstart[tid] = *reinterpret_cast<const unsigned short *>(cbytes + rs_counter_off<BITS>(tid, 0u));
    if (tid == 0) tail.tstart[R] = tile_n;      // the padding of the last tile ranks after every element
    __syncthreads();   // counters are dead: the same memory now stages the sorted tile
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) stage[pos[k]] = make_uint2(key[k], val[k]);
    if (tid <= G::MASK) {
      // this tile's elements of digit `tid`, published; then the sum over the tiles before it
      const uint32_t tstart = tail.tstart[tid], count = tail.tstart[tid + 1] - tstart;
      unsigned long long *mine = tile_state + (size_t)t * R + tid;
      uint32_t excl = 0;
      if (t == 0) st_relaxed_u64(mine, sweep_word(tag, SWEEP_FLAG_INC, count));
      else {
        st_relaxed_u64(mine, sweep_word(tag, SWEEP_FLAG_AGG, count));
        uint32_t p = t;                       // tiles [0, p) are still to be accounted for
        bool found = false;
        // the tile just before is examined alone first (in a long sort it usually holds its inclusive
        // prefix already, and seven more loads would be wasted traffic); after that, SWEEP_LOOKBACK
        // tiles per step with their loads in flight together
        for (;;) {
          const unsigned long long w0 = ld_relaxed_u64(tile_state + (size_t)(p - 1u) * R + tid);
          const uint32_t flag = (uint32_t)(w0 >> 32) & 3u;
          if ((uint32_t)(w0 >> 34) == tag && flag != 0) {
            excl += (uint32_t)w0;
            --p;
            found = flag == SWEEP_FLAG_INC;
            break;
          }
          if (backoff_ns) __nanosleep(backoff_ns);
        }
        while (!found) {
          unsigned long long w[SWEEP_LOOKBACK];
#pragma unroll
          for (int j = 0; j < SWEEP_LOOKBACK; ++j)
            w[j] = p > (uint32_t)j ? ld_relaxed_u64(tile_state + (size_t)(p - 1u - j) * R + tid) : 0ull;
          const uint32_t p_before = p;
#pragma unroll
          for (int j = 0; j < SWEEP_LOOKBACK; ++j) {
            if (found || p == 0) break;
            const uint32_t flag = (uint32_t)(w[j] >> 32) & 3u;
            if ((uint32_t)(w[j] >> 34) != tag || flag == 0) break;      // not published yet: read again from here
            excl += (uint32_t)w[j];
            --p;
            found = flag == SWEEP_FLAG_INC;
          }
          // nothing new: the tile before is still ranking.  Do not hammer L2 (and the issue slots of
          // whatever shares the SM) while it does
          if (!found && p == p_before && backoff_ns) __nanosleep(backoff_ns);
        }
        st_relaxed_u64(mine, sweep_word(tag, SWEEP_FLAG_INC, excl + count));
      }
      tail.out_base[tid] = dbase + excl - tstart;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
      const uint32_t q = k * RS_THREADS + tid;
      if (q < tile_n) {
        const uint2 kv = stage[q];
        const uint32_t dst = tail.out_base[(kv.x >> shift) & G::MASK] + q;
        keys_out[dst] = kv.x;
        vals_out[dst] = kv.y;
      }
    }
    __syncthreads();
  }
  // the last CTA of the pass to leave puts the control words back to zero for the next sort
  if (tid == 0) {
    __threadfence();
    if (atomicAdd(&ctl[RS_SWEEP_EXIT + pass], 1u) + 1u == gridDim.x) tail.ticket = 0xffffffffu;
  }
  __syncthreads();
  if (tail.ticket == 0xffffffffu) {
    for (uint32_t k = tid; k < 256u; k += RS_THREADS) ctl[RS_SWEEP_HIST + pass * 256 + k] = 0u;
    if (tid == 0) { ctl[RS_SWEEP_TICKET + pass] = 0u; ctl[RS_SWEEP_EXIT + pass] = 0u; }
  }
}

void sort_plan(int nbits, int &passes, int &bits_per_pass, int max_bits) {
  if (nbits < 1) nbits = 1;
  if (max_bits < 1 || max_bits > 8) max_bits = RS_MAX_BITS;
  passes = (nbits + max_bits - 1) / max_bits;
  bits_per_pass = (nbits + passes - 1) / passes;
}

template <int BITS>
static void launch_pass(SortBuffers &sb, int cur, bool identity, const uint32_t *d_n, int shift, uint32_t grid,
                        const uint32_t *key_range, cudaStream_t st) {
  rs_hist_kernel<BITS><<<grid, RS_THREADS, rs_counter_words(BITS) * 4, st>>>(sb.keys[cur], d_n, sb.capacity, shift,
                                                                            sb.block_hist, sb.max_tiles, key_range);
  rs_scan_kernel<<<1u << BITS, 256, 0, st>>>(sb.block_hist, sb.max_tiles, d_n, sb.capacity, sb.digit_totals, key_range,
                                             shift);
  rs_scatter_kernel<BITS><<<grid, RS_THREADS, rs_scatter_smem(BITS), st>>>(
      sb.keys[cur], identity ? nullptr : sb.vals[cur], sb.keys[cur ^ 1], sb.vals[cur ^ 1], d_n, sb.capacity, shift,
      sb.block_hist, sb.max_tiles, sb.digit_totals, key_range);
}

// pause of a look-back that found nothing new (ns; GSPLAT_SORT_BACKOFF_NS for experiments)
static uint32_t sweep_backoff_ns() {
  static const uint32_t ns = [] { const char *e = getenv("GSPLAT_SORT_BACKOFF_NS"); return e ? (uint32_t)atoi(e) : 100u; }();
  return ns;
}

template <int BITS>
static void launch_sweep(SortBuffers &sb, int cur, bool identity, const uint32_t *d_n, int shift, int pass, uint32_t grid,
                         const uint32_t *key_range, cudaStream_t st) {
  rs_sweep_kernel<BITS><<<grid, RS_THREADS, rs_scatter_smem(BITS), st>>>(
      sb.keys[cur], identity ? nullptr : sb.vals[cur], sb.keys[cur ^ 1], sb.vals[cur ^ 1], d_n, sb.capacity, shift, pass,
      sb.tile_state, sb.sweep_ctl, key_range, sweep_backoff_ns());
}

template <int BITS> static cudaError_t sort_attr() {
  cudaError_t e = cudaFuncSetAttribute(rs_scatter_kernel<BITS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)rs_scatter_smem(BITS));
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(rs_sweep_kernel<BITS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rs_scatter_smem(BITS));
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(rs_hist_kernel<BITS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              (int)rs_counter_words(BITS) * 4);
}
// Dynamic shared memory opt-in of every digit width on the CURRENT device (once per device).
cudaError_t setup_sort_kernels() {
  cudaError_t e = cudaSuccess;
  auto acc = [&](cudaError_t r) { if (e == cudaSuccess) e = r; };
  acc(sort_attr<1>()); acc(sort_attr<2>()); acc(sort_attr<3>()); acc(sort_attr<4>());
  acc(sort_attr<5>()); acc(sort_attr<6>()); acc(sort_attr<7>()); acc(sort_attr<8>());
  return e;
}

int launch_radix_sort(SortBuffers &sb, const uint32_t *d_n, uint32_t n_upper, int begin_bit, int end_bit,
                      bool identity_vals, cudaStream_t st, int max_bits, const uint32_t *key_range) {
  if (n_upper > sb.capacity) n_upper = sb.capacity;
  const uint32_t ntiles = (uint32_t)(((uint64_t)n_upper + RS_TILE - 1) / RS_TILE);
  int cur = 0;
  if (ntiles == 0) return cur;
  int passes, bpp;
  sort_plan(end_bit - begin_bit, passes, bpp, max_bits);
  const uint32_t grid = ntiles < (uint32_t)RS_GRID ? ntiles : (uint32_t)RS_GRID;
  if (sb.tile_state && sb.sweep_ctl && passes <= RS_SWEEP_MAX_PASSES) {
    // single-sweep form: the digit totals of all passes in one launch, then one launch per digit
    // (tiles are handed out by ticket, so any grid works: sb.sweep_grid caps it)
    const uint32_t grid_cap = sb.sweep_grid ? sb.sweep_grid : (uint32_t)RS_GRID;
    const uint32_t sgrid = grid < grid_cap ? grid : grid_cap;
    ++g_launches;
    // GSPLAT_SORT_SERIAL_WRAP (tests): wipe the look-back words every that-many sorts instead of every 2^26
    const uint32_t wrap_mask = [] {
      const char *e = getenv("GSPLAT_SORT_SERIAL_WRAP");
      uint32_t period = e ? (uint32_t)atoi(e) : 0u;
      if (period < 2u || (period & (period - 1u))) period = 1u << 26;      // a power of two
      return period - 1u;
    }();
    rs_hist_all_kernel<<<grid, RS_THREADS, 0, st>>>(sb.keys[0], d_n, sb.capacity, begin_bit, end_bit, bpp, passes,
                                                     sb.sweep_ctl, key_range, sb.tile_state,
                                                     (size_t)256 * sb.max_tiles, wrap_mask);
    for (int p = 0; p < passes; ++p) {
      const int shift = begin_bit + p * bpp;
      const int bits = (end_bit - shift) < bpp ? (end_bit - shift) : bpp;
      const bool ident = p == 0 && identity_vals;
      ++g_launches;
      switch (bits) {
        case 1: launch_sweep<1>(sb, cur, ident, d_n, shift, p, sgrid, key_range, st); break;
        case 2: launch_sweep<2>(sb, cur, ident, d_n, shift, p, sgrid, key_range, st); break;
        case 3: launch_sweep<3>(sb, cur, ident, d_n, shift, p, sgrid, key_range, st); break;
        case 4: launch_sweep<4>(sb, cur, ident, d_n, shift, p, sgrid, key_range, st); break;
        case 5: launch_sweep<5>(sb, cur, ident, d_n, shift, p, sgrid, key_range, st); break;
        case 6: launch_sweep<6>(sb, cur, ident, d_n, shift, p, sgrid, key_range, st); break;
        case 7: launch_sweep<7>(sb, cur, ident, d_n, shift, p, sgrid, key_range, st); break;
        default: launch_sweep<8>(sb, cur, ident, d_n, shift, p, sgrid < 148u ? sgrid : 148u, key_range, st); break;
      }
      cur ^= 1;
    }
    return cur;
  }
  for (int p = 0; p < passes; ++p) {
    const int shift = begin_bit + p * bpp;
    const int bits = (end_bit - shift) < bpp ? (end_bit - shift) : bpp;
    const bool ident = p == 0 && identity_vals;
    g_launches += 3;
    switch (bits) {
      case 1: launch_pass<1>(sb, cur, ident, d_n, shift, grid, key_range, st); break;
      case 2: launch_pass<2>(sb, cur, ident, d_n, shift, grid, key_range, st); break;
      case 3: launch_pass<3>(sb, cur, ident, d_n, shift, grid, key_range, st); break;
      case 4: launch_pass<4>(sb, cur, ident, d_n, shift, grid, key_range, st); break;
      case 5: launch_pass<5>(sb, cur, ident, d_n, shift, grid, key_range, st); break;
      case 6: launch_pass<6>(sb, cur, ident, d_n, shift, grid, key_range, st); break;
      case 7: launch_pass<7>(sb, cur, ident, d_n, shift, grid, key_range, st); break;
      default: launch_pass<8>(sb, cur, ident, d_n, shift, grid < 148u ? grid : 148u, key_range, st); break;
    }
    cur ^= 1;
  }
  return cur;
}

// ---- K4: per-tile ranges over the sorted tile keys (counts = end - start) -------
// Each thread inspects four consecutive keys (one 128-bit load) plus the key before them: a range
// starts where a key differs from its predecessor and ends where it differs from its successor.
__global__ void __launch_bounds__(256)
tile_ranges_kernel(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ d_n, uint32_t cap,
                   uint2 *__restrict__ ranges, uint32_t num_tiles) {
  const uint32_t n = num_elems(d_n, cap);
  const uint32_t nquads = (n + 3) / 4;
  for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < nquads; q += gridDim.x * blockDim.x) {
    const uint32_t i0 = q * 4;
    uint32_t k[4];
    if (i0 + 4 <= n) {
      const uint4 v = *reinterpret_cast<const uint4 *>(keys + i0);
      k[0] = v.x; k[1] = v.y; k[2] = v.z; k[3] = v.w;
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e) k[e] = i0 + e < n ? keys[i0 + e] : 0xffffffffu;
    }
    uint32_t prev = i0 ? keys[i0 - 1] : 0xffffffffu;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const uint32_t i = i0 + e;
      if (i < n && k[e] != prev) {
        if (k[e] < num_tiles) ranges[k[e]].x = i;
        if (i && prev < num_tiles) ranges[prev].y = i;
      }
      prev = k[e];
    }
    if (i0 + 4 >= n) {   // the thread holding the last element closes the last range
      const uint32_t last = k[(n - 1) - i0];
      if (last < num_tiles) ranges[last].y = n;
    }
  }
}

void launch_tile_ranges(const uint32_t *keys, const uint32_t *d_n, uint32_t n_upper, uint32_t cap, uint2 *ranges,
                        uint32_t num_tiles, cudaStream_t st) {
  if (!n_upper) return;   // `ranges` arrives zeroed (frame_reset)
  const int blocks = min(ceil_div(n_upper, 1024), 148 * 8);
  ++g_launches;
  tile_ranges_kernel<<<blocks, 256, 0, st>>>(keys, d_n, cap, ranges, num_tiles);
}

}  // namespace gsplat

// ---- K3 without any allocation: gsplat_sort_pairs_static ----------------------------------------
// No counterpart in the reference; added for callers that
// cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
// it).  Same three kernels per digit as launch_radix_sort; everything the sort needs besides the
// caller's input and output arrays — the ping-pong buffer of the passes in between, the per-tile
// digit histograms, the digit totals and the element count word — lives in ONE file-scope device
// array of fixed size.  Every GPU that loads the library holds it (8.5 MiB) until the process exits.
//
// The array is shared by every call on a device and keeps the last call's contents: the words that
// must start at zero are cleared by the call itself, and calls that overlap on the GPU (two streams,
// two host threads) RACE — serialise them.
#include "../../include/gsplat_b200.h"
#include <mutex>
#include <stdio.h>

namespace gsplat {
namespace {

constexpr uint32_t STATIC_SORT_MAX = GSPLAT_SORT_STATIC_MAX;                 // elements per call
static_assert(STATIC_SORT_MAX % RS_TILE == 0, "whole sort tiles");
constexpr uint32_t STATIC_SORT_TILES = STATIC_SORT_MAX / RS_TILE;            // 512
// layout of the scratch array, in 32-bit words (every part starts on a 256-byte boundary)
constexpr size_t SS_KEYS = 0;                                                // [MAX] intermediate keys
constexpr size_t SS_VALS = SS_KEYS + STATIC_SORT_MAX;                        // [MAX] intermediate values
constexpr size_t SS_HIST = SS_VALS + STATIC_SORT_MAX;                        // [256 x TILES] per-tile digit counts
constexpr size_t SS_TOTALS = SS_HIST + (size_t)256 * STATIC_SORT_TILES;      // [256] digit totals
constexpr size_t SS_COUNT = SS_TOTALS + 256;                                 // [64] word 0 = n
constexpr size_t SS_WORDS = SS_COUNT + 64;

}  // namespace

// In HOST code this name is a host-side stand-in, not a device address: handing it to a kernel
// faults (or, where the GPU can reach host memory, silently uses the wrong memory).  The host gets
// the device address once per device from cudaGetSymbolAddress (static_sort_scratch below).
static __device__ __align__(256) uint32_t g_static_sort_scratch[SS_WORDS];

// sets the element count and clears the digit totals (the only words of the scratch that are read
// before they are written when a digit's histogram rows are shorter than the last call's)
__global__ void static_sort_begin_kernel(uint32_t *scratch, uint32_t n) {
  const uint32_t t = threadIdx.x;
  if (t < 256) scratch[SS_TOTALS + t] = 0u;
  if (t < 64) scratch[SS_COUNT + t] = t == 0 ? n : 0u;
}

namespace {

// Device address of the scratch array on `device` (which must be current), resolved once per
// device; also the per-device dynamic-shared-memory opt-in of the sort kernels.
cudaError_t static_sort_scratch(int device, uint32_t **out) {
  static std::mutex mu;
  static uint32_t *addr[256] = {nullptr};
  if (device < 0 || device >= 256) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[device]) {
    uint32_t *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void **)&p, g_static_sort_scratch);
    if (e != cudaSuccess) return e;
    e = setup_sort_kernels();
    if (e != cudaSuccess) return e;
    addr[device] = p;
  }
  *out = addr[device];
  return cudaSuccess;
}

int static_fail(char *err, size_t err_len, int code, const char *what, const char *detail) {
  if (err && err_len) snprintf(err, err_len, "gsplat_sort_pairs_static: %s%s%s", what, detail ? ": " : "", detail ? detail : "");
  return code;
}

}  // namespace
}  // namespace gsplat

extern "C" int gsplat_sort_pairs_static(int device, const uint32_t *d_keys_in, const uint32_t *d_values_in,
                                        uint32_t *d_keys_out, uint32_t *d_values_out, size_t n, int begin_bit,
                                        int end_bit, void *stream, char *err, size_t err_len) {
  using namespace gsplat;
  if (err && err_len) err[0] = 0;
  if (begin_bit < 0 || end_bit > 32 || begin_bit >= end_bit)
    return static_fail(err, err_len, GSPLAT_E_ARG, "bit range must satisfy 0 <= begin_bit < end_bit <= 32", nullptr);
  if (n > STATIC_SORT_MAX) {
    char msg[96];
    snprintf(msg, sizeof msg, "n = %zu exceeds the static scratch (GSPLAT_SORT_STATIC_MAX = %u elements)", n,
             (unsigned)STATIC_SORT_MAX);
    return static_fail(err, err_len, GSPLAT_E_CAPACITY, msg, nullptr);
  }
  if (n == 0) return GSPLAT_OK;
  if (!d_keys_in || !d_values_in || !d_keys_out || !d_values_out)
    return static_fail(err, err_len, GSPLAT_E_ARG, "null device pointer", nullptr);
  if (((uintptr_t)d_keys_in | (uintptr_t)d_values_in | (uintptr_t)d_keys_out | (uintptr_t)d_values_out) & 15u)
    return static_fail(err, err_len, GSPLAT_E_ARG, "device pointers must be 16-byte aligned", nullptr);
  if (d_keys_in == d_keys_out || d_values_in == d_values_out)
    return static_fail(err, err_len, GSPLAT_E_ARG, "input and output arrays must be distinct", nullptr);
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return static_fail(err, err_len, GSPLAT_E_CUDA, "cudaSetDevice", cudaGetErrorString(e));
  uint32_t *scratch = nullptr;
  e = static_sort_scratch(device, &scratch);
  if (e != cudaSuccess) return static_fail(err, err_len, GSPLAT_E_CUDA, "cudaGetSymbolAddress / kernel set-up", cudaGetErrorString(e));
  cudaStream_t st = (cudaStream_t)stream;

  static_sort_begin_kernel<<<1, 256, 0, st>>>(scratch, (uint32_t)n);
  ++g_launches;
  const uint32_t *d_n = scratch + SS_COUNT;
  int passes, bpp;
  sort_plan(end_bit - begin_bit, passes, bpp, RS_MAX_BITS);
  const uint32_t ntiles = (uint32_t)((n + RS_TILE - 1) / RS_TILE);
  const uint32_t grid = ntiles < (uint32_t)RS_GRID ? ntiles : (uint32_t)RS_GRID;
  // pass p reads where pass p - 1 wrote (pass 0: the caller's input) and writes the caller's output
  // when an even number of passes follows it, the scratch otherwise: the last pass lands in the output
  const uint32_t *src_k = d_keys_in, *src_v = d_values_in;
  for (int p = 0; p < passes; ++p) {
    const bool to_out = ((passes - 1 - p) & 1) == 0;
    SortBuffers sb{};
    sb.keys[0] = const_cast<uint32_t *>(src_k); sb.vals[0] = const_cast<uint32_t *>(src_v);
    sb.keys[1] = to_out ? d_keys_out : scratch + SS_KEYS;
    sb.vals[1] = to_out ? d_values_out : scratch + SS_VALS;
    sb.block_hist = scratch + SS_HIST; sb.digit_totals = scratch + SS_TOTALS;
    sb.capacity = STATIC_SORT_MAX; sb.max_tiles = STATIC_SORT_TILES;
    const int shift = begin_bit + p * bpp;
    const int bits = (end_bit - shift) < bpp ? (end_bit - shift) : bpp;
    g_launches += 3;
    switch (bits) {
      case 1: launch_pass<1>(sb, 0, false, d_n, shift, grid, nullptr, st); break;
      case 2: launch_pass<2>(sb, 0, false, d_n, shift, grid, nullptr, st); break;
      case 3: launch_pass<3>(sb, 0, false, d_n, shift, grid, nullptr, st); break;
      case 4: launch_pass<4>(sb, 0, false, d_n, shift, grid, nullptr, st); break;
      case 5: launch_pass<5>(sb, 0, false, d_n, shift, grid, nullptr, st); break;
      case 6: launch_pass<6>(sb, 0, false, d_n, shift, grid, nullptr, st); break;
      default: launch_pass<7>(sb, 0, false, d_n, shift, grid, nullptr, st); break;
    }
    src_k = sb.keys[1]; src_v = sb.vals[1];
  }
  e = cudaGetLastError();
  if (e != cudaSuccess) return static_fail(err, err_len, GSPLAT_E_CUDA, "kernel launch", cudaGetErrorString(e));
  return GSPLAT_OK;
}
