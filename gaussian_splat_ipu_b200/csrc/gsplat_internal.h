// gsplat_internal.h — shared declarations of the sm_100a kernels behind libgsplat_b200.so.
// Stage names follow SURVEY.md §8: K1 project, (scan), K2 duplicate, K3 sort, K4 ranges, K5 blend.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

namespace gsplat {

// Per-frame constants, computed once on the host exactly as GSplat::renderInternal does
// (codelets/splat/codelets.cpp:443-448,470,622-628) and passed to K1 by value.
struct FrameParams {
  float mvp[16];     // projection * modelview, glm column-major: mvp[c*4+r]
  float tanfov;      // (float)tan(0.5*fov): used for BOTH axes (codelets.cpp:444-445)
  float focal_x, focal_y;
  float scale_div;   // fxy[1]
  float img_w, img_h;
  float tile_w, tile_h;
  float guard;       // |tile diagonal| * clipSize
  float guard_sq;    // smallest float q with sqrtf(q) >= guard: sqrtf(x) < guard  <=>  x < guard_sq (exactly)
  float inv_tile_w;  // 1 / tile_w when tile_w is a power of two (x / tile_w == x * inv_tile_w exactly), else 0
  int tiles_across, tiles_down;
  int row_begin, row_end;   // tile-row band rendered by this context
};

// Scene in HBM: the 56 live bytes of the reference's 256-byte-strided record plus
// the camera-independent part of the projection, cached.  w == 1 and gid == index+1 are implicit;
// empty slots have a NaN position.
// Gaussian3D::ComputeCov3D (ipu_geometry.hpp:315-323) depends on the Gaussian and on fxy[1] (the
// scale divisor) only, not on the camera: its nine f32 results are computed once per (scene,
// divisor) by cov3_kernel — same operations, same order, so the same bits as the reference's
// per-frame evaluation — and K1 reads them (36 B) instead of quaternion + log-scales (28 B) and
// the three expf, the normalisation and two 3x3 products per Gaussian per frame.
struct SceneSoA {
  float4 *pos_c;     // x, y, z, cov3[2][2]                      (K1 input)
  float4 *cov_a;     // cov3[0][0], [0][1], [0][2], [1][0]       (K1 input; glm [col][row])
  float4 *cov_b;     // cov3[1][1], [1][2], [2][0], [2][1]       (K1 input)
  float4 *rgba;      // r, g, b, opacity (raw)                    (K1 input, drawn Gaussians)
  float4 *quat;      // w, x, y, z   (rot_0 = w, ipu_geometry.hpp:317)   (cov3_kernel input)
  float4 *scl;       // log-scale.x, .y, .z, unused                     (cov3_kernel input)
  uint32_t n;
};

// K1 output per Gaussian: one 48-byte AoS record (three float4, contiguous: one 48-byte bulk copy
// moves it into K5's shared-memory ring), written for drawn Gaussians only, plus a 32-bit depth
// key for EVERY Gaussian (culled = 0xffffffff, sorts last).
//   rec[3g+0] = mean.x, mean.y, hx, hy      half extents of the box around the ellipse alpha >= 1/255
//                                           (-inf: never visible; +inf: no finite extent, always visited)
//   rec[3g+1] = conic.x, conic.y, conic.z, qmax   Gaussian2D::ComputeConicOpacity's values; a pair whose
//                                           quadratic form exceeds qmax has alpha < 1/255 (skipped)
//   rec[3g+2] = opacity (0 when det == 0), r, g, b
struct Projected {
  float4 *rec;
  uint32_t *depth_key;   // ~bits(clip z) for drawn Gaussians (z < 0  =>  key < 2^31), per Gaussian
  uint2 *rect;           // tile rectangle {c0 | c1<<16, r0 | r1<<16}, rows clamped to this context's
                         // band; RECT_EMPTY when not drawn.  Read by the counts, the selection and K2
};
// ---- chained scan with decoupled look-back (K2 and the slice selection) ------------------------
// One 64-bit word per chain link: flag in the top two bits, value below, so one store publishes both.
constexpr unsigned long long LB_AGGREGATE = 1ull << 62;   // value = this link's own total
constexpr unsigned long long LB_PREFIX = 2ull << 62;      // value = inclusive prefix up to this link
constexpr unsigned long long LB_VALUE = (1ull << 62) - 1;
#ifndef GSPLAT_LB_WINDOW
#define GSPLAT_LB_WINDOW 1
#endif
constexpr int LB_WINDOW = GSPLAT_LB_WINDOW;   // predecessors inspected per lane and round (32 * LB_WINDOW links);
                                              // measured (profiles/micro/chain.cu): 128-link rounds are slower than 32
#ifdef __CUDACC__
__device__ __forceinline__ unsigned long long lb_load(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void lb_store(unsigned long long *p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// Called by ONE full warp of link `b` after it knows its own `total`: publishes it, walks back over
// the predecessors until one with a full prefix is found, publishes this link's inclusive prefix
// and returns the exclusive one (all lanes).  Every predecessor is already running (links take
// their position from an atomic ticket), so the spin on an unpublished word terminates.
__device__ __forceinline__ unsigned long long lb_exclusive_prefix(unsigned long long *lookback, uint32_t b,
                                                                  unsigned long long total, uint32_t lane) {
  unsigned long long excl = 0;
  if (b == 0) {
    if (lane == 0) lb_store(&lookback[0], LB_PREFIX | total);
    return 0;
  }
  if (lane == 0) lb_store(&lookback[b], LB_AGGREGATE | total);
  int pos = (int)b - 1;
  while (true) {
    unsigned long long w[LB_WINDOW];
#pragma unroll
    for (int j = 0; j < LB_WINDOW; ++j) {   // distance from b-1: lane + 32 j; issue all loads, then wait
      const int q = pos - (int)lane - 32 * j;
      w[j] = q >= 0 ? lb_load(&lookback[q]) : LB_PREFIX;   // before the chain start: prefix 0
    }
#pragma unroll
    for (int j = 0; j < LB_WINDOW; ++j) {
      const int q = pos - (int)lane - 32 * j;
      while ((w[j] >> 62) == 0) w[j] = lb_load(&lookback[q]);
    }
    bool found = false;
    unsigned long long v = 0;
#pragma unroll
    for (int j = 0; j < LB_WINDOW; ++j) {
      const uint32_t is_prefix = __ballot_sync(0xffffffffu, (w[j] >> 62) == 2);
      if (!found) {
        const int first = is_prefix ? __ffs(is_prefix) - 1 : 32;   // nearest link of this row with a full prefix
        if ((int)lane <= first) v += w[j] & LB_VALUE;
        found = is_prefix != 0;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    excl += v;
    if (found) break;
    pos -= 32 * LB_WINDOW;
  }
  if (lane == 0) lb_store(&lookback[b], LB_PREFIX | (excl + total));
  return excl;
}
#endif

#ifdef __CUDACC__
// expf as the host libm computes it.  The reference's codelets call exp() on floats
// (ipu_geometry.hpp:319-321, codelets.cpp:400); host-compiled, that is glibc's expf
// (sysdeps/ieee754/flt-32/e_expf.c, glibc >= 2.27; the algorithm of ARM's optimized-routines):
// with N = 32,
//   k = round(x * N / ln2),  r = x * N / ln2 - k,  exp(x) = 2^(k/N) * P(r),
// 2^(k/N) from a 32-entry table of doubles (exponent bits added by integer arithmetic), P a cubic,
// all in f64, rounded once to f32.  Restated here with the same table, constants and operation
// order, so the result is bit-identical to the oracle's (a correctly rounded exp differs from it
// for about one argument in 2000) at a fraction of the f64 work of exp(double).
// On x86-64 glibc dispatches expf through an ifunc: CPUs with FMA + AVX2 (every host this runs
// beside) get __expf_fma, the same source compiled with contraction — five of its products are
// fused into the following add (disassembly of libm 2.39: vfmadd132sd for z + SHIFT, vfmsub132sd
// for r, vfmadd213sd / vfmadd132sd x 2 for the polynomial).  GSPLAT_EXPF_FMA = 1 (default) restates
// that variant (8 f64 operations), 0 the unfused one (12).  Both agree with the host's expf on every
// one of 3e8 (CPU) / 6.7e7 (GPU, profiles/micro/exp_check.cu) sampled arguments in [-16, 0.8].
#ifndef GSPLAT_EXPF_FMA
#define GSPLAT_EXPF_FMA 1
#endif
static __constant__ unsigned long long EXPF_TAB[32] = {
    0x3ff0000000000000ull, 0x3fefd9b0d3158574ull, 0x3fefb5586cf9890full, 0x3fef9301d0125b51ull,
    0x3fef72b83c7d517bull, 0x3fef54873168b9aaull, 0x3fef387a6e756238ull, 0x3fef1e9df51fdee1ull,
    0x3fef06fe0a31b715ull, 0x3feef1a7373aa9cbull, 0x3feedea64c123422ull, 0x3feece086061892dull,
    0x3feebfdad5362a27ull, 0x3feeb42b569d4f82ull, 0x3feeab07dd485429ull, 0x3feea47eb03a5585ull,
    0x3feea09e667f3bcdull, 0x3fee9f75e8ec5f74ull, 0x3feea11473eb0187ull, 0x3feea589994cce13ull,
    0x3feeace5422aa0dbull, 0x3feeb737b0cdc5e5ull, 0x3feec49182a3f090ull, 0x3feed503b23e255dull,
    0x3feee89f995ad3adull, 0x3feeff76f2fb5e47ull, 0x3fef199bdd85529cull, 0x3fef3720dcef9069ull,
    0x3fef5818dcfba487ull, 0x3fef7c97337b9b5full, 0x3fefa4afa2a490daull, 0x3fefd0765b6e4540ull,
};
// table entry k of EXPF_TAB's shared-memory copy; `tab` may be a generic pointer or (TabRef) a 32-bit
// shared-window address, which spares the generic-to-shared conversion in an inner loop
struct TabRef { uint32_t saddr; };
__device__ __forceinline__ unsigned long long exp_tab_entry(const unsigned long long *tab, unsigned long long k) { return tab[k]; }
__device__ __forceinline__ unsigned long long exp_tab_entry(TabRef tab, unsigned long long k) {
  unsigned long long v;
  asm("ld.shared.u64 %0, [%1];" : "=l"(v) : "r"(tab.saddr + (uint32_t)k * 8u));
  return v;
}
// exp_ref_core's f64 constants.  exp_consts(true) reads them from constant memory, to be called ONCE
// outside a hot loop: the values then stay in registers (or become constant-bank operands) across
// the loop, where literal constants are re-materialised on every trip (8 moves per expf in K5's
// inner loop, profiles/) — ptxas folds anything whose value it can see.
struct ExpConsts { double invln2n, c0, c1, c2; };
static __constant__ double EXPF_K[4] = {0x1.71547652b82fep+0 * 32.0, 0x1.c6af84b912394p-5 / 32.0 / 32.0 / 32.0,
                                        0x1.ebfce50fac4f3p-3 / 32.0 / 32.0, 0x1.62e42ff0c52d6p-1 / 32.0};
__device__ __forceinline__ ExpConsts exp_consts(bool from_memory) {
  if (from_memory) return ExpConsts{EXPF_K[0], EXPF_K[1], EXPF_K[2], EXPF_K[3]};
  return ExpConsts{0x1.71547652b82fep+0 * 32.0, 0x1.c6af84b912394p-5 / 32.0 / 32.0 / 32.0,
                   0x1.ebfce50fac4f3p-3 / 32.0 / 32.0, 0x1.62e42ff0c52d6p-1 / 32.0};
}
// the main path of expf: valid for -103.97 < x < 88.72 (no NaN / overflow / underflow handling)
template <typename Tab>
__device__ __forceinline__ float exp_ref_core(float x, Tab tab /* EXPF_TAB in shared memory */, const ExpConsts &k) {
  constexpr double SHIFT = 0x1.8p+52;
  const double xd = (double)x;
#if GSPLAT_EXPF_FMA
  double kd = __fma_rn(k.invln2n, xd, SHIFT);
  const unsigned long long ki = (unsigned long long)__double_as_longlong(kd);
  kd = __dsub_rn(kd, SHIFT);
  const double r = __fma_rn(k.invln2n, xd, -kd);
  const double s = __longlong_as_double((long long)(exp_tab_entry(tab, ki & 31ull) + (ki << 47)));
  const double zz = __fma_rn(k.c0, r, k.c1), r2 = __dmul_rn(r, r);
  double y = __fma_rn(k.c2, r, 1.0);
  y = __fma_rn(zz, r2, y);
  return (float)__dmul_rn(y, s);
#else
  const double z = __dmul_rn(k.invln2n, xd);
  double kd = __dadd_rn(z, SHIFT);
  const unsigned long long ki = (unsigned long long)__double_as_longlong(kd);
  kd = __dsub_rn(kd, SHIFT);
  const double r = __dsub_rn(z, kd);
  const double s = __longlong_as_double((long long)(exp_tab_entry(tab, ki & 31ull) + (ki << 47)));
  const double zz = __dadd_rn(__dmul_rn(k.c0, r), k.c1), r2 = __dmul_rn(r, r);
  double y = __dadd_rn(__dmul_rn(k.c2, r), 1.0);
  y = __dadd_rn(__dmul_rn(zz, r2), y);
  return (float)__dmul_rn(y, s);
#endif
}
template <typename Tab>
__device__ __forceinline__ float exp_ref_core(float x, Tab tab) { return exp_ref_core(x, tab, exp_consts(false)); }
__device__ __forceinline__ float exp_ref(float x, const unsigned long long *tab /* EXPF_TAB in shared memory */) {
  if (!(x == x)) return x + x;
  if (x > 88.72283172607421875f) return __int_as_float(0x7f800000);    // 0x1.62e42ep6f: overflow
  if (x < -103.97207641601562500f) return 0.0f;                        // -0x1.9fe368p6f: underflow
  return exp_ref_core(x, tab);
}
#endif

constexpr uint32_t DEPTH_KEY_CULLED = 0xffffffffu;
constexpr uint32_t RECT_EMPTY = 1u;   // {lo = 1, hi = 0}: the rectangle of a Gaussian that is not drawn

constexpr int K1_THREADS = 256;
// counting re-run of K5 (gsplat_profile_tiles): words [0] visited pairs, [1] blended pairs, then
// K5_PROBE_WORDS experiment counters, then per tile {cycles, visited pairs}
constexpr int K5_PROBE_WORDS = 8;
constexpr int K5_TILE_WORDS0 = 2 + K5_PROBE_WORDS;
// K1 (and the parity inject) write the frame's band {row_begin, row_end} at this word offset from
// their drawn counter; tile_order and K5 read it there (`band` arguments below)
constexpr int BAND_WORD_OFFSET = 2;

// ---- launchers (all asynchronous on `st`) ------------------------------------------------
void launch_aos_to_soa(const float *aos16, SceneSoA scene, cudaStream_t st);
// the cached ComputeCov3D results of every Gaussian for scale divisor fxy[1] = scale_div
void launch_cov3(const SceneSoA &scene, float scale_div, cudaStream_t st);
// K1: writes Projected; drawn_counter += number of drawn Gaussians (V).
void launch_project(const SceneSoA &scene, const FrameParams &fp, Projected out, uint32_t *drawn_counter,
                    float *dbg_rec8, int32_t *dbg_rect4, bool dense, cudaStream_t st);
// Is `func` one of K1's kernels (to find its node in a captured frame graph and refresh its camera argument)?
bool is_project_kernel(const void *func);
// Debug/parity entry: fill depth key / rect from caller-supplied projected inputs.
void launch_inject(const float *depth, const int32_t *rect4, const uint8_t *drawn, uint32_t n,
                   const FrameParams &fp, Projected out, uint32_t *drawn_counter, cudaStream_t st);

// K3: stable LSD radix sort of (u32 key, u32 value) pairs over key bits [begin_bit, end_bit), in
// ceil(bits/7) digits of equal width.  n is read on the device from *d_n (clamped to capacity).
struct SortBuffers {
  uint32_t *keys[2];
  uint32_t *vals[2];
  uint32_t *block_hist;     // 256 * max_tiles
  uint32_t *digit_totals;   // 256
  uint32_t capacity;        // elements
  uint32_t max_tiles;
  // single-sweep passes (launch_radix_sort, RS_SWEEP): per (tile, digit) look-back words and the
  // self-cleaning control block (RsSweepCtl); nullptr selects the three-kernel passes
  unsigned long long *tile_state;   // 256 * max_tiles
  uint32_t *sweep_ctl;              // RS_SWEEP_CTL_WORDS, zero at rest
  uint32_t sweep_grid;              // CTAs of a single-sweep launch (0 = RS_GRID); tiles are handed out by ticket
};
// sweep_ctl layout (words): digit totals of every pass [RS_SWEEP_MAX_PASSES][256], then per pass a
// tile ticket and an exit counter, then the sort serial number
constexpr int RS_SWEEP_MAX_PASSES = 8;
constexpr int RS_SWEEP_HIST = 0;
constexpr int RS_SWEEP_TICKET = RS_SWEEP_MAX_PASSES * 256;
constexpr int RS_SWEEP_EXIT = RS_SWEEP_TICKET + RS_SWEEP_MAX_PASSES;
constexpr int RS_SWEEP_SERIAL = RS_SWEEP_EXIT + RS_SWEEP_MAX_PASSES;
constexpr int RS_SWEEP_CTL_WORDS = RS_SWEEP_SERIAL + 8;
constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = 8;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;
constexpr int RS_MAX_BITS = 7;      // digit width limit of the counter-based ranking (128 digits)
constexpr int RS_GRID = 148 * 3;   // CTAs walk RS_TILE-element tiles with a grid stride
void sort_plan(int nbits, int &passes, int &bits_per_pass, int max_bits = RS_MAX_BITS);
// identity_vals: the first digit takes value = element index instead of reading vals[0].
// returns index (0/1) of the buffer holding the sorted result
// key_range (optional, device): {max key, max ~key} of the keys, as select.cu accumulates them.
// Digits above the highest bit in which the keys differ are the same for every key, so their
// passes are skipped ON THE DEVICE (the launches stay: the frame is a fixed graph).  Skipped passes
// are always the last ones; the result then sits in buffer rs_result_buffer(...), which the
// consumer evaluates on the device too (the host's return value assumes no pass was skipped).
int launch_radix_sort(SortBuffers &sb, const uint32_t *d_n, uint32_t n_upper, int begin_bit, int end_bit,
                      bool identity_vals, cudaStream_t st, int max_bits = RS_MAX_BITS,
                      const uint32_t *key_range = nullptr);
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t key_range_bits(const uint32_t *key_range) {
  const uint32_t mx = key_range[0], mn = ~key_range[1];
  return mx > mn ? 32u - (uint32_t)__clz(mx ^ mn) : 0u;   // no key, or all keys equal: nothing to sort
}
// Index (0/1) of the buffer holding the result of a sort of bits [0, passes * bits_per_pass) that
// started in buffer 0 and skipped the passes above the key range.
__device__ __forceinline__ uint32_t rs_result_buffer(const uint32_t *key_range, uint32_t bits_per_pass, uint32_t passes) {
  const uint32_t need = (key_range_bits(key_range) + bits_per_pass - 1u) / bits_per_pass;
  return (need < passes ? need : passes) & 1u;
}
#endif

// Depth-slice selection (select.cu).  tau[s] = depth-key quantile at fractions f[s] of the drawn
// Gaussians (sampled); launch_select compacts, in index order, the (key, index) pairs of the
// Gaussians with *tau_lo <= key < *tau_hi (nullptr = open end) whose rectangle covers a tile that
// is still being blended (tile_done != nullptr), and writes their number to *total_out.
// `lookback` (ceil(n / 4096) + 1 words) and *ticket must be zero on entry.
struct SliceFractions {
  int count;
  float f[7];
};
void launch_depth_quantiles(const uint32_t *depth_key, uint32_t n, const SliceFractions &fr, uint32_t *tau,
                            cudaStream_t st);
void launch_select(const Projected &p, uint32_t n, const uint32_t *tau_lo, const uint32_t *tau_hi, int tiles_across,
                   const uint32_t *live_sat, uint32_t *keys_out, uint32_t *idx_out, unsigned long long *lookback,
                   uint32_t *ticket, uint32_t *total_out, uint32_t *key_range /* += {max key, max ~key} */,
                   cudaStream_t st);

// K2: walks the *d_count (<= count_cap) depth-sorted Gaussians of one slice (sorted_gauss = values
// of sort A) and emits one (tile id, Gaussian index) pair per tile of each Gaussian's rectangle —
// skipping tiles whose tile_done flag is set when tile_done != nullptr.  Offsets come from a
// single-pass chained scan (decoupled look-back over `lookback`, which must be zeroed — together
// with *ticket — before each launch), so the output is in (depth, gid, tile) order and
// bit-reproducible.  Writes the number of pairs to *total; sets *overflow (and writes nothing past
// the buffer) if it exceeds capacity.
constexpr int K2_THREADS = 256;
constexpr int K2_ITEMS = 4;
constexpr int K2_LOG_CHUNK = 10;
constexpr int K2_CHUNK = K2_THREADS * K2_ITEMS;   // Gaussians per CTA == per chain link
static_assert((1 << K2_LOG_CHUNK) == K2_CHUNK, "binary search depth");
// sorted_gauss[2] + key_range + the depth sort's plan: the buffer the (device-side pass skipping)
// sort left its result in is chosen on the device; key_range == nullptr: sorted_gauss[0].
void launch_duplicate(const Projected &p, const uint32_t *const sorted_gauss[2], const uint32_t *key_range,
                      uint32_t sort_bits_per_pass, uint32_t sort_passes, const uint32_t *d_count, uint32_t count_cap,
                      int tiles_across, const uint8_t *tile_done, const uint32_t *live_sat, uint32_t *keys, uint32_t *vals,
                      uint32_t capacity, unsigned long long *lookback, uint32_t *ticket, uint32_t *total,
                      uint32_t *overflow, cudaStream_t st);
// Summed-area table over the tiles whose tile_done flag is clear ((tiles_down+1) x (tiles_across+1)
// words) and the query used by the later depth slices: live tiles inside a tile rectangle.
void launch_live_sat(const uint8_t *tile_done, int tiles_across, int tiles_down, uint32_t *sat, cudaStream_t st);
__device__ __forceinline__ uint32_t live_tiles_in_rect(const uint32_t *__restrict__ sat, int tiles_across, uint32_t c0,
                                                       uint32_t c1, uint32_t r0, uint32_t r1) {
  const uint32_t gw = (uint32_t)tiles_across + 1u;
  return (__ldg(&sat[(r1 + 1) * gw + c1 + 1]) - __ldg(&sat[r0 * gw + c1 + 1])) -
         (__ldg(&sat[(r1 + 1) * gw + c0]) - __ldg(&sat[r0 * gw + c0]));
}
// a11: counts[tile] = drawn Gaussians whose rectangle covers the tile (the reference's list length).
// `diff` is a (tiles_down+1) x (tiles_across+1) int grid that must be zero on entry and is left zero.
constexpr size_t TILE_COUNT_MAX_SMEM = 200 * 1024;
size_t tile_count_smem(int tiles_across, int tiles_down);
void launch_tile_counts(const Projected &p, uint32_t n, int tiles_across, int tiles_down, int *diff, uint32_t *counts,
                        cudaStream_t st);

// K4: per-tile [start,end) over the sorted tile keys.
void launch_tile_ranges(const uint32_t *keys, const uint32_t *d_n, uint32_t n_upper, uint32_t cap,
                        uint2 *ranges, uint32_t num_tiles, cudaStream_t st);

// K5: front-to-back blend of every tile of the band over one depth slice's lists; `first`/`last`
// say whether pixel state is initialised / saved (tstate: T per pixel, -1 = stopped; tile_done: all
// pixels of the tile stopped).  pair_counters != nullptr selects the counting variant (roofline
// accounting): [0] += visited pairs, [1] += blended pairs, then per tile {cycles, visited pairs}.
// `split`: CTAs per semantic tile (each blends tile_h / split rows; must divide tile_h).
// `order`: scratch of one word per tile; launch_blend fills it with the band's tiles, longest list
// first, and K5 takes its tiles in that order.
struct TileState {
  uint8_t *tile_done;    // [tiles] every pixel of the tile has stopped (read by K2 of later slices)
  uint8_t *part_done;    // [tiles * 4 (of BLEND_MAX_SPLIT slots)] every pixel of the tile's 8-column strip has stopped
  uint32_t *parts_done;  // [tiles] strips of the tile that have stopped (4 = the whole tile)
};
constexpr int BLEND_MAX_SPLIT = 8;
inline size_t tile_state_bytes(int tiles) { return (size_t)tiles * (1 + BLEND_MAX_SPLIT + 4) + 16; }
// `wide`: the 53-register form of K5 (a context that renders one frame at a time; blend.cu).
// `band`: device words {row_begin, row_end} of the frame's tile-row band (written by K1); the grid
// covers every tile of the image and the CTAs beyond the band leave at once.
void launch_blend(const Projected &p, const uint32_t *sorted_vals, const uint2 *ranges, const FrameParams &fp,
                  const uint32_t *band, float4 *fb_tiled, float *tstate, const TileState &ts, int split, bool wide,
                  uint32_t *order, bool first, bool last, unsigned long long *pair_counters, cudaStream_t st);

// a10: tiled f32 RGBA -> de-tiled BGR8 rows.
void launch_bgr8(const float4 *fb_tiled, uint8_t *bgr, int W, int H, int tile_w, int tile_h,
                 cudaStream_t st);
void launch_bgr8_rows(const float4 *fb_tiled, uint8_t *bgr, int W, int y0, int y1, int tile_w, int tile_h,
                      cudaStream_t st);

// cudaFuncAttributeMaxDynamicSharedMemorySize is a property of (kernel, device context): every
// device a context is created on needs its own opt-in.  gsplat_create calls these once per device
// (mutex-guarded table in gsplat_api.cu), with that device current.
cudaError_t setup_blend_kernels();
cudaError_t setup_duplicate_kernels();
cudaError_t setup_sort_kernels();

extern unsigned long long g_launches;   // kernels launched by this library (bench accounting)
inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

}  // namespace gsplat
