// blend.cu — K5: front-to-back alpha compositing of each semantic tile's z-ordered list, and the
// a10 framebuffer post-process.
//
// Follows GSplat::renderTile (codelets/splat/codelets.cpp:358-421) and
// Gaussian2D::ComputeConicOpacity (include/splat/ipu_geometry.hpp:278-286):
//   per pixel (integer coordinates as floats, no +0.5), T = 1, C = 0; for each Gaussian of the
//   tile's list in z order:  d = mean - pixel;  power = -0.5*(A dx^2 + C dy^2) - B dx dy;
//   power > 0 -> skip;  alpha = min(0.99, opacity * exp(power));  alpha < 1/255 -> skip;
//   test_T = T (1 - alpha);  test_T < 1e-4 -> stop this pixel (the stopping Gaussian is not added);
//   C += colour * alpha * T over all FOUR channels (4th = opacity);  T = test_T.
// There is no per-pixel bbox test in the reference: a Gaussian contributes to exactly the pixels
// of the 32x20 tiles its rectangle covers, so the CTA *is* the semantic tile (or a column strip).
//
// BIT-EXACT blend chain.  The three tests above are discontinuities: a last-bit difference in
// alpha moves T, and T * (1 - alpha) < 1e-4 then flips for pixels that sit on the threshold —
// stacks of opaque splats (alpha capped at 0.99, T = 1 -> 0.01 -> 1e-4) sit there systematically,
// and one flip changes a pixel by up to colour * 0.99 * 0.01, beyond the 1/255 parity bar
// (measured on the shipped PLY orbit: 10 of 256 frames with a fused / ex2.approx evaluation).
// Every value that feeds a decision or the accumulators is therefore computed with the
// reference's operations in the reference's order: separately rounded f32 products and sums (this
// file is compiled with -fmad=false) and exp() as the host libm computes it (exp_ref,
// gsplat_internal.h).  The framebuffer is bit-identical to the host-compiled reference's.
//
// What makes that affordable is a conservative, FUSED pre-test: q = d^T Q d evaluated with two FMAs
// against the per-Gaussian bound qmax (K1) rejects the pairs whose alpha is certainly < 1/255 —
// the vast majority — and only the surviving lanes run the exact chain.
//
// Mapping: one CTA per part of a tile (a column strip), one CONSUMER warp per 8 x 4 pixel block
// with one pixel per thread in registers, and one PRODUCER warp.  The producer streams the tile's
// list through a shared-memory ring: per list entry ONE 48-byte bulk copy (cp.async.bulk, the TMA
// unit; SASS UBLKCP) of the K1 record, completion counted in bytes by an mbarrier per ring stage.
// Consumer warps wait on a stage's "full" barrier, test the staged entries against their block
// (lane = entry: box of the alpha >= 1/255 ellipse vs the block, one ballot), walk the hits, and
// release the stage through its "empty" barrier.  No CTA-wide barrier inside the loop: a warp
// whose block few splats reach runs ahead of its neighbours by up to the ring depth.
// FP32-pipe bound; no tensor cores (each pair has its own quadratic form).
#include "gsplat_internal.h"
#include <stdlib.h>

namespace gsplat {

unsigned long long g_launches = 0;

constexpr float ALPHA_MIN = 1.0f / 255.0f;

#ifndef GSPLAT_BLEND_STAGES
#define GSPLAT_BLEND_STAGES 3
#endif
#ifndef GSPLAT_BLEND_BATCH
#define GSPLAT_BLEND_BATCH 128
#endif
#ifndef GSPLAT_BLEND_AHEAD
#define GSPLAT_BLEND_AHEAD 2
#endif
constexpr int BLEND_STAGES = GSPLAT_BLEND_STAGES;    // ring depth
constexpr uint32_t BLEND_AHEAD = GSPLAT_BLEND_AHEAD; // stages between an entry's index load and its record copy
constexpr int BLEND_BATCH = GSPLAT_BLEND_BATCH;      // list entries per ring stage (multiple of 32)
static_assert(BLEND_BATCH % 32 == 0, "consumers walk a stage in chunks of 32 entries");
constexpr int BLEND_MAX_WARPS = 20;                  // consumer warps of an undivided 32 x 20 tile
// How the producer warp moves a record into the ring:
//   1  cp.async.bulk (TMA unit, SASS UBLKCP): one 48-byte copy per entry, bytes counted by the
//      stage's mbarrier.  UBLKCP takes its addresses from UNIFORM registers, so a warp issues its 32
//      copies one lane at a time (ELECT / R2UR loop, 8 instructions per entry).
//   0  cp.async (SASS LDGSTS): three 16-byte copies per entry with per-lane addresses, the lane's
//      completion delivered to the stage's mbarrier by cp.async.mbarrier.arrive.noinc.
// Both are kept (-DGSPLAT_BLEND_TMA=0/1); DESIGN.md §4 has the measurement that picked the default.
#ifndef GSPLAT_BLEND_TMA
#define GSPLAT_BLEND_TMA 0
#endif
// Measured and dropped (profiles/ab_k5_r2.txt): the entry-by-entry inner loop (each surviving entry
// runs the exact chain for the lanes it reaches: 4 % slower than the per-pixel sets below) and 64
// staged entries per consumer turn (one phase B over two chunks: K5 0.373-0.379 ms against 0.348-0.350).
// 1 = phase A walks a compacted list of the hit lanes, 0 = find-first-set on the hit ballot.
// Measured (profiles/ab_k5_r2c.txt): the list is SLOWER, K5 0.331 ms against 0.314 (the store, the warp
// barrier and the dependent load cost more than the uniform-datapath bit walk): default 0.
#ifndef GSPLAT_BLEND_HITLIST
#define GSPLAT_BLEND_HITLIST 0
#endif
// 1 = a consumer warp tests all of a stage's entries against its block before it walks any (0 = chunk by chunk)
#ifndef GSPLAT_BLEND_STAGETEST
#define GSPLAT_BLEND_STAGETEST 1
#endif
// 1 = phase B's turn is straight-line (selects), 0 = branches per test (measured: 0.331 against 0.361 ms)
#ifndef GSPLAT_BLEND_PRED
#define GSPLAT_BLEND_PRED 1
#endif
// Measured and dropped (profiles/ab_k5_r2e.txt): phase B with TWO entries per turn (their alphas evaluated side
// by side for instruction-level parallelism, applied in list order): K5 0.322 against 0.313 ms.
// 1 = keep expf's f64 constants and the table address in registers across the blend loop (0.331 against 0.341 ms)
#ifndef GSPLAT_BLEND_HOIST
#define GSPLAT_BLEND_HOIST 1
#endif
constexpr uint32_t BLEND_FULL_ARRIVALS = GSPLAT_BLEND_TMA ? 1u : 32u;   // arrivals that complete a "full" phase

struct BlendSmem {
  float4 ring[BLEND_STAGES][BLEND_BATCH * 3];        // 48 B per entry: rec[3g .. 3g+2]
  unsigned long long full[BLEND_STAGES];             // mbarrier: the stage's bytes have landed
  unsigned long long empty[BLEND_STAGES];            // mbarrier: every consumer warp is done with the stage
  unsigned long long exp_tab[32];                    // exp_ref's 2^(k/32) table
#if GSPLAT_BLEND_HITLIST
  uint32_t hits[BLEND_MAX_WARPS][34];                // per consumer warp: lanes of the chunk's hit entries, in order
#endif
  uint32_t warps_done;                               // consumer warps whose 32 pixels have all stopped
  uint32_t strip_done[4];                            // the same per 8-pixel column strip of the part
  uint32_t stop;                                     // producer: nothing more will be staged
};

// ---- mbarrier / bulk-copy primitives (PTX ISA 8.x, sm_90+) ---------------------------------------
__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
// Blocks until the phase of `bar` with the given parity has completed.
//   GSPLAT_BLEND_WAIT_NS = 0: try_wait with a suspend-time hint (GSPLAT_BLEND_WAIT_HINT ns): the thread
//     is suspended in hardware until the phase completes or the hint runs out.
//   GSPLAT_BLEND_WAIT_NS > 0: plain try_wait, then nanosleep of that many ns before the next try: a
//     waiting warp issues a handful of instructions per sleep instead of re-polling (measured in the
//     round-2 capture: 13 % of K5's warp instructions were re-polls of the hinted form, ~90 per wait).
#ifndef GSPLAT_BLEND_WAIT_NS
#define GSPLAT_BLEND_WAIT_NS 0
#endif
#ifndef GSPLAT_BLEND_WAIT_HINT
#define GSPLAT_BLEND_WAIT_HINT 0x989680
#endif
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, uint32_t parity) {
#if GSPLAT_BLEND_WAIT_NS > 0
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "WAIT_%=:\n"
      "nanosleep.u32 %2;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@!p bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_addr(bar)),
      "r"(parity), "r"((uint32_t)GSPLAT_BLEND_WAIT_NS)
      : "memory");
#else
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_addr(bar)),
      "r"(parity), "r"((uint32_t)GSPLAT_BLEND_WAIT_HINT)
      : "memory");
#endif
}
// 48-byte global -> shared bulk copy, completion signalled on `bar` in bytes (complete_tx)
__device__ __forceinline__ void bulk_copy_48(void *dst, const void *src, unsigned long long *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 48, [%2];" ::"r"(smem_addr(dst)),
               "l"(src), "r"(smem_addr(bar))
               : "memory");
}
// 16-byte global -> shared asynchronous copy (L2 only: the record is not re-read by this SM)
__device__ __forceinline__ void async_copy_16(void *dst, const void *src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr(dst)), "l"(src) : "memory");
}
// the executing thread's earlier cp.async copies arrive on `bar` when they have completed; the
// pending count is NOT incremented: the arrival is one of the barrier's expected arrivals
__device__ __forceinline__ void async_copies_arrive(unsigned long long *bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ uint32_t ld_volatile_shared(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(v) : "r"(smem_addr(p)) : "memory");
  return v;
}

// One list entry against one pixel, in the reference's arithmetic (this TU is -fmad=false: every
// * and + below is its own IEEE operation, as the host build of the codelet evaluates them).
// codelets.cpp:389-410; returns true when the pixel stops at this entry.
template <bool COUNT>
__device__ __forceinline__ bool blend_exact(const float4 b, const float4 c, float dx, float dy,
                                            const TabRef exp_tab, const ExpConsts &ek, float &T, float &Cr,
                                            float &Cg, float &Cb, float &Ca, unsigned long long &n_blend) {
  const float power = -0.5f * (b.x * dx * dx + b.z * dy * dy) - b.y * dx * dy;
  if (power > 0.0f) return false;
  // survivors of the pre-test have -13 < power <= 0: exp_ref's NaN / overflow / underflow branches
  // cannot be taken and are skipped
  const float e = c.x * exp_ref_core(power, exp_tab, ek);
  const float alpha = e < 0.99f ? e : 0.99f;        // glm::min(0.99f, e)
  if (alpha < ALPHA_MIN) return false;
  const float test_T = T * (1.f - alpha);
  if (test_T < 0.0001f) return true;
  Cr = Cr + c.y * alpha * T;                        // (colour * alpha) * T, then the accumulate
  Cg = Cg + c.z * alpha * T;
  Cb = Cb + c.w * alpha * T;
  Ca = Ca + c.x * alpha * T;                        // 4th channel: the opacity itself (colour.w)
  T = test_T;
  if (COUNT) ++n_blend;
  return false;
}

// The same step as straight-line code: `on` false leaves the pixel untouched whatever the entry
// holds (a lane without work is handed a stale ring entry).  Same operations on the same values in
// the same order for the lanes that are on, so the results are those of blend_exact bit for bit.
template <bool COUNT>
__device__ __forceinline__ bool blend_exact_pred(bool on, const float4 b, const float4 c, float dx, float dy,
                                                 const TabRef exp_tab, const ExpConsts &ek, float &T, float &Cr,
                                                 float &Cg, float &Cb, float &Ca, unsigned long long &n_blend) {
  const float power = -0.5f * (b.x * dx * dx + b.z * dy * dy) - b.y * dx * dy;
  const float e = c.x * exp_ref_core(power, exp_tab, ek);
  const float alpha = e < 0.99f ? e : 0.99f;
  const float test_T = T * (1.f - alpha);
  const bool ok = on && !(power > 0.0f) && !(alpha < ALPHA_MIN);
  const bool stop = ok && test_T < 0.0001f;
  const bool upd = ok && !stop;
  const float ar = c.y * alpha * T, ag = c.z * alpha * T, ab = c.w * alpha * T, aa = c.x * alpha * T;
  Cr = upd ? Cr + ar : Cr;
  Cg = upd ? Cg + ag : Cg;
  Cb = upd ? Cb + ab : Cb;
  Ca = upd ? Ca + aa : Ca;
  T = upd ? test_T : T;
  if (COUNT) n_blend += upd ? 1ull : 0ull;
  return stop;
}

// Depth slices (see duplicate.cu): the list of a tile may arrive in several consecutive pieces, one
// launch each.  Between launches a pixel's state lives in the framebuffer (C) and in `tstate`
// (T, or -1 once the pixel has stopped); a tile whose pixels have all stopped raises
// tile_done[tile] so that later slices neither emit duplicates for it nor launch work on it.
// FIRST: start from T = 1, C = 0.  LAST: no state is saved.
// blockDim = (consumer warps + 1) * 32; the last warp is the producer.
template <bool COUNT, bool FIRST, bool LAST>
__device__ __forceinline__ void
blend_body(const Projected &p, const uint32_t *__restrict__ vals, const uint2 *__restrict__ ranges,
           int tiles_across, const uint32_t *__restrict__ band, int full_tile_h, int split_log2, float4 *__restrict__ fb,
           float *__restrict__ tstate, const TileState &ts, const uint32_t *__restrict__ order,
           unsigned long long *__restrict__ pair_counters) {
  extern __shared__ __align__(128) unsigned char blend_smem_raw[];
  BlendSmem &sm = *reinterpret_cast<BlendSmem *>(blend_smem_raw);
  const long long t_begin = COUNT ? clock64() : 0;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t NW = (blockDim.x >> 5) - 1;           // consumer warps
  // a CTA blends a "part" of a semantic tile: all its rows, `part_w` = 32 >> split_log2 of its columns
  // (the whole tile, or a column strip so that a crowded tile's list is walked by several SMs at
  // once); parts of one tile are neighbours in the grid.  A warp owns an 8 x 4 pixel block: a splat
  // a few pixels across then occupies a third of the lanes of the warps it reaches, where a
  // 32 x 1 pixel row would leave nine lanes in ten idle.  (Shifts and masks only: this prologue is
  // run by every warp of every CTA, two thirds of which find an empty list on the bench frame.)
  const uint32_t sl = (uint32_t)split_log2, part_w = 32u >> sl, sub = blockIdx.x & ((1u << sl) - 1u);
  const uint32_t slot = blockIdx.x >> sl;
  // the grid covers the whole image; this context's band (device words written by K1) may be less
  const uint32_t row_begin = band[0], row_end = band[1];
  if (slot >= (row_end - row_begin) * (uint32_t)tiles_across) return;
  // tiles are taken longest list first (tile_order_kernel, which hands out (row, column) of the
  // band's tiles packed in one word) so the long ones do not start last
  const uint32_t packed = order[slot];
  const uint32_t tcol = packed & 0xffffu, trow = row_begin + (packed >> 16);
  const uint32_t tile = trow * (uint32_t)tiles_across + tcol;
  const uint2 range = ranges[tile];
  const uint32_t n = range.y - range.x;
  // "done" flags are kept per 8-pixel column STRIP of the tile (4 per tile) so that slices may cut
  // the tile into different numbers of parts (the later, emptier slices use narrower parts)
  const uint32_t strips = part_w >> 3, strip0 = tile * 4u + sub * strips;
  // consumer geometry: this warp's 8 x 4 block and this lane's pixel
  const uint32_t cb = warp & (strips - 1u), rb = warp >> (2u - sl);
  const uint32_t x_local = sub * part_w + cb * 8u + (lane & 7u), y_local = rb * 4u + (lane >> 3);
  const bool valid = y_local < (uint32_t)full_tile_h;    // tile heights that are no multiple of 4
  // a10 layout: tile-major, row-major inside the tile
  const size_t px = (size_t)tile * (32 * full_tile_h) + (size_t)(valid ? y_local : 0u) * 32 + x_local;
  if (n == 0) {
    // nothing (new) in this part's list.  First slice: the pixels' initial state is the result.
    if (FIRST && warp != NW && valid) {
      fb[px] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (!LAST) tstate[px] = 1.f;
    }
    return;
  }
  if (!FIRST) {
    bool all_done = true;                                 // every pixel of the part has stopped: its state stands
    for (uint32_t k = 0; k < strips; ++k) all_done = all_done && ts.part_done[strip0 + k];
    if (all_done) return;
  }

  if (tid == 0) {
    for (int s = 0; s < BLEND_STAGES; ++s) { mbar_init(&sm.full[s], BLEND_FULL_ARRIVALS); mbar_init(&sm.empty[s], NW); }
    sm.warps_done = 0;
    sm.strip_done[0] = sm.strip_done[1] = sm.strip_done[2] = sm.strip_done[3] = 0;
    sm.stop = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (tid < 32) sm.exp_tab[tid] = EXPF_TAB[tid];
  __syncthreads();

  const uint32_t nbatch = (n + BLEND_BATCH - 1) / BLEND_BATCH;

  if (warp == NW) {
    // ---- producer: the tile's list, BLEND_BATCH entries per stage --------------------------------
    // The list indices are loaded BLEND_AHEAD stages before their records are copied: with the ring
    // not full (consumers that reject nearly everything: the long lists of the later slices), a stage
    // costs the producer 1 / BLEND_AHEAD of an index round trip instead of a whole one, and the record
    // copies of up to BLEND_STAGES stages are in flight behind it.
    constexpr uint32_t PER_LANE = BLEND_BATCH / 32;
    uint32_t gidx[BLEND_AHEAD][PER_LANE];
#pragma unroll
    for (uint32_t a = 0; a < BLEND_AHEAD; ++a)
#pragma unroll
      for (uint32_t q = 0; q < PER_LANE; ++q) {
        const uint32_t j = a * BLEND_BATCH + q * 32 + lane;
        gidx[a][q] = j < n ? __ldg(&vals[range.x + j]) : 0u;
      }
    for (uint32_t it = 0; it < nbatch; ++it) {
      const uint32_t stage = it % BLEND_STAGES, use = it / BLEND_STAGES;
      mbar_wait(&sm.empty[stage], (use & 1u) ^ 1u);        // first use of a stage: passes at once
      if (ld_volatile_shared(&sm.warps_done) == NW) {      // every pixel of the part has stopped
        if (lane == 0) *reinterpret_cast<volatile uint32_t *>(&sm.stop) = 1u;
        __syncwarp();
        if (lane < BLEND_FULL_ARRIVALS) mbar_arrive(&sm.full[stage]);   // releases the consumers waiting on this stage
        break;
      }
      const uint32_t base = it * BLEND_BATCH, cnt = min((uint32_t)BLEND_BATCH, n - base);
#if GSPLAT_BLEND_TMA
      if (lane == 0) mbar_arrive_expect_tx(&sm.full[stage], cnt * 48u);
      __syncwarp();
#endif
#pragma unroll
      for (uint32_t q = 0; q < PER_LANE; ++q) {
        const uint32_t j = q * 32 + lane;
        if (j < cnt) {
          const float4 *src = p.rec + (size_t)gidx[0][q] * 3;
          float4 *dst = &sm.ring[stage][j * 3];
#if GSPLAT_BLEND_TMA
          bulk_copy_48(dst, src, &sm.full[stage]);
#else
          async_copy_16(dst, src); async_copy_16(dst + 1, src + 1); async_copy_16(dst + 2, src + 2);
#endif
        }
      }
#if !GSPLAT_BLEND_TMA
      async_copies_arrive(&sm.full[stage]);   // every lane: one of the 32 arrivals, once its copies have landed
#endif
#pragma unroll
      for (uint32_t q = 0; q < PER_LANE; ++q) {            // indices of stage it + BLEND_AHEAD
#pragma unroll
        for (uint32_t a = 0; a + 1 < BLEND_AHEAD; ++a) gidx[a][q] = gidx[a + 1][q];
        const uint32_t j = base + BLEND_AHEAD * BLEND_BATCH + q * 32 + lane;
        gidx[BLEND_AHEAD - 1][q] = j < n ? __ldg(&vals[range.x + j]) : 0u;
      }
    }
  } else {
    // ---- consumers ---------------------------------------------------------------------------
    const float xf = (float)(tcol * 32u + x_local);
    const float yf = (float)(trow * full_tile_h) + (float)y_local;
    // centre of the warp's 8 x 4 block (pixels x0..x0+7, y0..y0+3)
    const float bcx = (float)(tcol * 32u + sub * part_w + cb * 8u) + 3.5f;
    const float bcy = (float)(trow * full_tile_h + rb * 4u) + 1.5f;

#if GSPLAT_BLEND_HOIST
    // loop invariants the compiler would otherwise rebuild on every trip of the blend loop
    uint32_t exp_tab_addr = smem_addr(sm.exp_tab);
    asm volatile("" : "+r"(exp_tab_addr));
    const TabRef exp_tab{exp_tab_addr};
    const ExpConsts ek = exp_consts(true);
#else
    const TabRef exp_tab{smem_addr(sm.exp_tab)};
    const ExpConsts ek = exp_consts(false);
#endif
#if GSPLAT_BLEND_HITLIST
    uint32_t *my_hits = sm.hits[warp];
    uint32_t lanemask_lt;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(lanemask_lt));
#endif
    float T = 1.f, Cr = 0.f, Cg = 0.f, Cb = 0.f, Ca = 0.f;
    bool done = !valid;
    if (!FIRST && valid) {
      const float4 c = fb[px];
      Cr = c.x; Cg = c.y; Cb = c.z; Ca = c.w;
      T = tstate[px];
      done = T < 0.f;
    }
    uint32_t stop_at = done ? 0u : n;   // COUNT: list entries this pixel visits (reference semantics)
    unsigned long long n_blend = 0;     // COUNT
    bool warp_done = __all_sync(0xffffffffu, done);
    if (warp_done && lane == 0) { atomicAdd(&sm.strip_done[cb], 1u); atomicAdd(&sm.warps_done, 1u); }

    for (uint32_t it = 0; it < nbatch; ++it) {
      const uint32_t stage = it % BLEND_STAGES, use = it / BLEND_STAGES;
      mbar_wait(&sm.full[stage], use & 1u);
      if (ld_volatile_shared(&sm.stop)) break;
      if (!warp_done) {
        const uint32_t base = it * BLEND_BATCH, cnt = min((uint32_t)BLEND_BATCH, n - base);
        const float4 *ring = sm.ring[stage];
        // lane = entry: does the box around its alpha >= 1/255 ellipse reach this warp's block?
        // The WHOLE stage is tested first (BLEND_BATCH / 32 independent loads and ballots per lane): a
        // warp's instructions issue ~16 cycles apart, so testing the chunks one after the other made
        // a stage without a single hit cost four dependent test sequences — and the long, mostly
        // irrelevant lists of the later slices are walked at exactly that latency (they were the
        // critical path of those launches: SMs active 70 % of the launch, profiles/).
        constexpr uint32_t CHUNKS = BLEND_BATCH / 32;
        uint32_t masks[CHUNKS];
#if GSPLAT_BLEND_STAGETEST
#pragma unroll
        for (uint32_t h = 0; h < CHUNKS; ++h) {
          const uint32_t jj = h * 32 + lane;
          bool hit = false;
          if (jj < cnt) {
            const float4 a = ring[jj * 3];
            hit = fabsf(a.x - bcx) <= a.z + 3.5f && fabsf(a.y - bcy) <= a.w + 1.5f;
          }
          masks[h] = __ballot_sync(0xffffffffu, hit);
        }
#endif
#pragma unroll
        for (uint32_t h = 0; h < CHUNKS; ++h) {
          const uint32_t c0 = h * 32;
#if GSPLAT_BLEND_STAGETEST
          uint32_t mask = masks[h];
#else
          if (c0 >= cnt) break;
          const uint32_t jj = c0 + lane;
          bool hit = false;
          if (jj < cnt) {
            const float4 a = ring[jj * 3];
            hit = fabsf(a.x - bcx) <= a.z + 3.5f && fabsf(a.y - bcy) <= a.w + 1.5f;
          }
          uint32_t mask = __ballot_sync(0xffffffffu, hit);
#endif
          if (mask == 0) continue;
          // Phase A, whole warp per entry: the fused pre-test q = A dx^2 + 2 B dx dy + C dy^2 against
          // qmax (alpha < 1/255 beyond it) gives every pixel the set of chunk entries it must blend.
          // Two entries per turn (their loads and pre-tests overlap).
          uint32_t smask = 0;
#if GSPLAT_BLEND_HITLIST
          // the hit lanes write their lane numbers, in order, to the warp's list: the walk below is an
          // indexed loop (one 8-byte load per two entries) instead of find-first-set on the ballot
          const uint32_t nh = __popc(mask);
          if (hit) {
            const uint32_t pos = __popc(mask & lanemask_lt);
            my_hits[pos] = lane;
            if (pos + 1 == nh) my_hits[nh] = lane;          // an odd count walks its last entry twice
          }
          __syncwarp();
          for (uint32_t i = 0; i < nh; i += 2) {
            const uint2 ll = *reinterpret_cast<const uint2 *>(my_hits + i);
            const uint32_t l0 = ll.x, l1 = ll.y;
#else
          while (mask) {
            const uint32_t l0 = __ffs(mask) - 1;
            mask &= mask - 1;
            const uint32_t l1 = mask ? __ffs(mask) - 1 : l0;
            mask &= mask - 1;
#endif
            const float4 a0 = ring[(c0 + l0) * 3], b0 = ring[(c0 + l0) * 3 + 1];
            const float4 a1 = ring[(c0 + l1) * 3], b1 = ring[(c0 + l1) * 3 + 1];
            const float dx0 = a0.x - xf, dy0 = a0.y - yf, dx1 = a1.x - xf, dy1 = a1.y - yf;
            const float q0 = fmaf(b0.z * dy0, dy0, dx0 * fmaf(b0.x, dx0, (b0.y + b0.y) * dy0));
            const float q1 = fmaf(b1.z * dy1, dy1, dx1 * fmaf(b1.x, dx1, (b1.y + b1.y) * dy1));
            smask |= (q0 <= b0.w ? 1u << l0 : 0u) | (q1 <= b1.w ? 1u << l1 : 0u);
          }
          if (done) smask = 0;
          if (COUNT) {   // probes: chunks with a hit, hit entries, hit entries no live pixel passes the pre-test for
            uint32_t hm = masks[h], dead = 0;
            while (hm) { const uint32_t l = __ffs(hm) - 1; hm &= hm - 1; dead += __any_sync(0xffffffffu, (smask >> l) & 1u) ? 0u : 1u; }
            const uint32_t live = __popc(__ballot_sync(0xffffffffu, !done));
            if (lane == 0) {
              atomicAdd(&pair_counters[2], 1ull); atomicAdd(&pair_counters[3], (unsigned long long)__popc(masks[h])); atomicAdd(&pair_counters[4], (unsigned long long)dead);
              if (live == 32) { atomicAdd(&pair_counters[7], (unsigned long long)__popc(masks[h])); atomicAdd(&pair_counters[8], (unsigned long long)dead); }
              atomicAdd(&pair_counters[9], (unsigned long long)live);
            }
          }
          // Phase B, one pixel per lane: every lane walks ITS OWN entries in list order through the
          // exact chain, so a turn of the loop blends up to 32 different (pixel, entry) pairs instead of
          // one entry against the few pixels it reaches; turns = the longest per-pixel set.
          while (__any_sync(0xffffffffu, smask != 0)) {
            if (COUNT) {   // probes: phase B turns and the lanes busy in them
              const uint32_t busy = __popc(__ballot_sync(0xffffffffu, smask != 0));
              if (lane == 0) { atomicAdd(&pair_counters[5], 1ull); atomicAdd(&pair_counters[6], (unsigned long long)busy); }
            }
#if GSPLAT_BLEND_PRED
            // straight-line turn: lanes whose set is empty run a stale entry with `on` false.  With
            // ~15 lanes busy per turn (measured) a warp-level skip almost never happens, and the
            // branches + reconvergence points cost more issue slots than they save.
            const bool on = smask != 0;
            const uint32_t kl = (uint32_t)(__ffs(smask) - 1) & 31u;
            smask &= smask - 1;
            const float4 *e = ring + (c0 + kl) * 3;
            const float2 mxy = *reinterpret_cast<const float2 *>(e);
            if (blend_exact_pred<COUNT>(on, e[1], e[2], mxy.x - xf, mxy.y - yf, exp_tab, ek, T, Cr, Cg, Cb, Ca, n_blend)) {
              done = true;
              smask = 0;
              if (COUNT) stop_at = base + c0 + kl + 1;
            }
#else
            if (smask) {
              const uint32_t kl = __ffs(smask) - 1;
              smask &= smask - 1;
              const float4 *e = ring + (c0 + kl) * 3;
              const float2 mxy = *reinterpret_cast<const float2 *>(e);
              if (blend_exact<COUNT>(e[1], e[2], mxy.x - xf, mxy.y - yf, exp_tab, ek, T, Cr, Cg, Cb, Ca, n_blend)) {
                done = true;
                smask = 0;
                if (COUNT) stop_at = base + c0 + kl + 1;
              }
            }
#endif
          }
          if (__all_sync(0xffffffffu, done)) { warp_done = true; break; }
        }
        if (warp_done && lane == 0) { atomicAdd(&sm.strip_done[cb], 1u); atomicAdd(&sm.warps_done, 1u); }   // before the release below
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.empty[stage]);
    }

    if (!LAST && valid) tstate[px] = done ? -1.f : T;
    if (COUNT) {
      unsigned long long n_visit = stop_at;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        n_visit += __shfl_xor_sync(0xffffffffu, n_visit, o);
        n_blend += __shfl_xor_sync(0xffffffffu, n_blend, o);
      }
      if (lane == 0) {
        atomicAdd(&pair_counters[0], n_visit); atomicAdd(&pair_counters[1], n_blend);
        atomicAdd(&pair_counters[K5_TILE_WORDS0 + 2 * (size_t)tile + 1], n_visit);            // per-tile profile tap
        atomicAdd(&pair_counters[K5_TILE_WORDS0 + 2 * (size_t)tile], (unsigned long long)(clock64() - t_begin) / NW);
      }
    }
    if (valid) fb[px] = make_float4(Cr, Cg, Cb, Ca);
  }
  if (!LAST) {
    __syncthreads();   // every consumer warp has added itself to the counters if its pixels have stopped
    const uint32_t row_blocks = NW >> (2u - sl);
    if (tid < strips && sm.strip_done[tid] == row_blocks && !ts.part_done[strip0 + tid]) {
      ts.part_done[strip0 + tid] = 1;
      if (atomicAdd(&ts.parts_done[tile], 1u) + 1u == 4u) ts.tile_done[tile] = 1;   // last strip of the tile to stop
    }
  }
}

// The two entry points of the body.  blend_kernel: any CTA size up to a whole tile (21 warps), 40
// registers.  blend_kernel_wide: CTAs of at most 11 warps (two or four strips per tile) with the 53
// registers the body asks for, 3 CTAs per SM instead of 4.  Measured (profiles/ab_k5_r2e.txt): the wide
// form is the faster K5 when a frame runs alone (0.292 against 0.311 ms) and the slower one when
// several frames are in flight (1578 against 1608 frames/s: residency counts there), so a context that
// renders one frame at a time launches the wide form and the lanes of a dispatcher the other.
// (Capping at 32 registers, 5 CTAs per SM: K5 alone 0.341 ms, frames in flight +0.7 %: not taken.)
constexpr int BLEND_WIDE_WARPS = 11;
template <bool COUNT, bool FIRST, bool LAST>
__global__ void __launch_bounds__((BLEND_MAX_WARPS + 1) * 32)
blend_kernel(Projected p, const uint32_t *__restrict__ vals, const uint2 *__restrict__ ranges,
             int tiles_across, const uint32_t *__restrict__ band, int full_tile_h, int split_log2, float4 *__restrict__ fb,
             float *__restrict__ tstate, TileState ts, const uint32_t *__restrict__ order,
             unsigned long long *__restrict__ pair_counters) {
  blend_body<COUNT, FIRST, LAST>(p, vals, ranges, tiles_across, band, full_tile_h, split_log2, fb, tstate, ts, order, pair_counters);
}
template <bool FIRST, bool LAST>
__global__ void __launch_bounds__(BLEND_WIDE_WARPS * 32, 3)
blend_kernel_wide(Projected p, const uint32_t *__restrict__ vals, const uint2 *__restrict__ ranges,
                  int tiles_across, const uint32_t *__restrict__ band, int full_tile_h, int split_log2, float4 *__restrict__ fb,
                  float *__restrict__ tstate, TileState ts, const uint32_t *__restrict__ order) {
  blend_body<false, FIRST, LAST>(p, vals, ranges, tiles_across, band, full_tile_h, split_log2, fb, tstate, ts, order, nullptr);
}

// ---- launch order of the tiles: longest list first ---------------------------------------------
// CTAs are dispatched in blockIdx order; a tile with a long list that starts late leaves the other
// SMs idle at the end of the launch.  One CTA buckets the band's tiles by list length (power-of-two
// class + 2 mantissa bits, longest class first; tiles that already stopped count as empty) and
// writes the permutation K5 indexes with blockIdx, each tile as (row within the band) << 16 | column
// so that K5's prologue needs no division.  Order inside a class is arbitrary: it only changes
// scheduling, never results.
constexpr int ORDER_BUCKETS = 33 * 4;
__device__ __forceinline__ uint32_t order_bucket(uint32_t len) {
  if (len == 0) return ORDER_BUCKETS - 1;
  const uint32_t e = 31u - (uint32_t)__clz(len);                       // floor(log2 len)
  const uint32_t m = e >= 2 ? (len >> (e - 2)) & 3u : (len << (2 - e)) & 3u;
  return (31u - e) * 4u + (3u - m);                                     // longer lists -> smaller bucket
}
__global__ void __launch_bounds__(1024)
tile_order_kernel(const uint2 *__restrict__ ranges, int tiles_across, const uint32_t *__restrict__ band,
                  const uint8_t *__restrict__ tile_done, uint32_t *__restrict__ order) {
  __shared__ uint32_t cnt[ORDER_BUCKETS];
  const uint32_t row_begin = band[0];
  const int ntiles = (int)((band[1] - row_begin) * (uint32_t)tiles_across);
  for (int k = threadIdx.x; k < ORDER_BUCKETS; k += blockDim.x) cnt[k] = 0;
  __syncthreads();
  const uint32_t t0 = (uint32_t)row_begin * tiles_across;
  for (int i = threadIdx.x; i < ntiles; i += blockDim.x) {
    const uint2 r = ranges[t0 + i];
    const uint32_t len = (tile_done && tile_done[t0 + i]) ? 0u : r.y - r.x;
    atomicAdd(&cnt[order_bucket(len)], 1u);
  }
  __syncthreads();
  if (threadIdx.x < 32) {   // exclusive scan of the bucket counts: one warp, ORDER_BUCKETS / 32 (+1) per lane
    constexpr int PER = (ORDER_BUCKETS + 31) / 32;
    uint32_t v[PER], sum = 0;
#pragma unroll
    for (int q = 0; q < PER; ++q) {
      const int k = threadIdx.x * PER + q;
      v[q] = k < ORDER_BUCKETS ? cnt[k] : 0u;
      sum += v[q];
    }
    uint32_t inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
      if (threadIdx.x >= o) inc += t;
    }
    uint32_t run = inc - sum;
#pragma unroll
    for (int q = 0; q < PER; ++q) {
      const int k = threadIdx.x * PER + q;
      if (k < ORDER_BUCKETS) cnt[k] = run;
      run += v[q];
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < ntiles; i += blockDim.x) {
    const uint2 r = ranges[t0 + i];
    const uint32_t len = (tile_done && tile_done[t0 + i]) ? 0u : r.y - r.x;
    order[atomicAdd(&cnt[order_bucket(len)], 1u)] = ((uint32_t)(i / tiles_across) << 16) | (uint32_t)(i % tiles_across);
  }
}

static void launch_tile_order(const uint2 *ranges, const FrameParams &fp, const uint32_t *band, const uint8_t *tile_done,
                              bool first, uint32_t *order, cudaStream_t st) {
  ++g_launches;
  tile_order_kernel<<<1, 1024, 0, st>>>(ranges, fp.tiles_across, band, first ? nullptr : tile_done, order);
}

// Dynamic shared memory per K5 CTA.  GSPLAT_K5_SMEM_PAD (bytes, experiments) pads it to lower the number
// of K5 CTAs resident per SM, which leaves room for another lane's K1 / sort CTAs beside them.
static size_t blend_smem_bytes() {
  static const size_t pad = [] { const char *e = getenv("GSPLAT_K5_SMEM_PAD"); return e ? (size_t)atol(e) : (size_t)0; }();
  return sizeof(BlendSmem) + pad;
}

template <bool COUNT, bool FIRST, bool LAST>
static void launch_blend_t(const Projected &p, const uint32_t *sorted_vals, const uint2 *ranges, const FrameParams &fp,
                           const uint32_t *band, float4 *fb_tiled, float *tstate, const TileState &ts, int split, bool wide,
                           const uint32_t *order, unsigned long long *pair_counters, cudaStream_t st) {
  const int rows = fp.tiles_down;   // launch geometry for the whole image: the band is only known on the device
  const int full_h = (int)fp.tile_h;
  const int split_log2 = split >= 4 ? 2 : (split >= 2 ? 1 : 0);  // split is 1, 2 or 4
  const int part_w = 32 >> split_log2;                            // columns per CTA
  const int warps = (part_w / 8) * ((full_h + 3) / 4);           // one consumer warp per 8 x 4 pixel block
  const unsigned grid = (unsigned)(rows * fp.tiles_across) << split_log2;
  ++g_launches;
  if constexpr (!COUNT) {
    if (wide && warps + 1 <= BLEND_WIDE_WARPS) {
      blend_kernel_wide<FIRST, LAST><<<grid, (warps + 1) * 32, blend_smem_bytes(), st>>>(
          p, sorted_vals, ranges, fp.tiles_across, band, full_h, split_log2, fb_tiled, tstate, ts, order);
      return;
    }
  }
  blend_kernel<COUNT, FIRST, LAST><<<grid, (warps + 1) * 32, blend_smem_bytes(), st>>>(
      p, sorted_vals, ranges, fp.tiles_across, band, full_h, split_log2, fb_tiled, tstate, ts, order, pair_counters);
}

template <bool COUNT, bool FIRST, bool LAST> static cudaError_t blend_attr() {
  cudaError_t e = cudaFuncSetAttribute(blend_kernel<COUNT, FIRST, LAST>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)blend_smem_bytes());
  if constexpr (!COUNT) {
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(blend_kernel_wide<FIRST, LAST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)blend_smem_bytes());
  }
  return e;
}
// Dynamic shared memory opt-in of every K5 variant on the CURRENT device (the attribute belongs to
// the device's context: gsplat_create runs this once per device).
cudaError_t setup_blend_kernels() {
  cudaError_t e = cudaSuccess;
  auto acc = [&](cudaError_t r) { if (e == cudaSuccess) e = r; };
  acc(blend_attr<false, false, false>()); acc(blend_attr<false, false, true>());
  acc(blend_attr<false, true, false>()); acc(blend_attr<false, true, true>());
  acc(blend_attr<true, false, false>()); acc(blend_attr<true, false, true>());
  acc(blend_attr<true, true, false>()); acc(blend_attr<true, true, true>());
  return e;
}

void launch_blend(const Projected &p, const uint32_t *sorted_vals, const uint2 *ranges, const FrameParams &fp,
                  const uint32_t *band, float4 *fb_tiled, float *tstate, const TileState &ts, int split, bool wide,
                  uint32_t *order, bool first, bool last, unsigned long long *pair_counters, cudaStream_t st) {
  launch_tile_order(ranges, fp, band, ts.tile_done, first, order, st);
#define GSPLAT_BLEND(C, F, L) launch_blend_t<C, F, L>(p, sorted_vals, ranges, fp, band, fb_tiled, tstate, ts, split, wide, order, pair_counters, st)
  if (pair_counters) {
    if (first) { if (last) GSPLAT_BLEND(true, true, true); else GSPLAT_BLEND(true, true, false); }
    else { if (last) GSPLAT_BLEND(true, false, true); else GSPLAT_BLEND(true, false, false); }
  } else {
    if (first) { if (last) GSPLAT_BLEND(false, true, true); else GSPLAT_BLEND(false, true, false); }
    else { if (last) GSPLAT_BLEND(false, false, true); else GSPLAT_BLEND(false, false, false); }
  }
#undef GSPLAT_BLEND
}

// ---- a10: IpuSplatter::getFrameBuffer, src/splat/ipu_rasteriser.cpp:115-144 -----------------
// min(v*255, 255) -> u8 (cv::saturate_cast: round half to even, clamp to [0,255]) -> RGBA2BGR ->
// de-tile.  One thread converts 4 horizontally adjacent pixels: 64 B in, 12 B out.
__device__ __forceinline__ uint32_t to_u8(float v) {
  v = v * 255.0f;
  v = v < 255.0f ? v : 255.0f;
  int i = __float2int_rn(v);
  i = i < 0 ? 0 : (i > 255 ? 255 : i);
  return (uint32_t)i;
}

__global__ void __launch_bounds__(256)
bgr8_kernel(const float4 *__restrict__ fb, uint8_t *__restrict__ bgr, int W, int y0, int y1, int tile_w, int tile_h,
            int tiles_across) {
  const int quads_per_row = W / 4;
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= quads_per_row * (y1 - y0)) return;
  const int y = y0 + q / quads_per_row, x = (q % quads_per_row) * 4;
  const int tile = (y / tile_h) * tiles_across + x / tile_w;
  const float4 *src = fb + (size_t)tile * (tile_w * tile_h) + (y % tile_h) * tile_w + (x % tile_w);
  uint32_t b[12];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float4 v = src[k];
    b[k * 3 + 0] = to_u8(v.z); b[k * 3 + 1] = to_u8(v.y); b[k * 3 + 2] = to_u8(v.x);
  }
  uint32_t *out = reinterpret_cast<uint32_t *>(bgr + ((size_t)y * W + x) * 3);
#pragma unroll
  for (int k = 0; k < 3; ++k)
    out[k] = b[k * 4] | (b[k * 4 + 1] << 8) | (b[k * 4 + 2] << 16) | (b[k * 4 + 3] << 24);
}

void launch_bgr8(const float4 *fb_tiled, uint8_t *bgr, int W, int H, int tile_w, int tile_h, cudaStream_t st) {
  launch_bgr8_rows(fb_tiled, bgr, W, 0, H, tile_w, tile_h, st);
}

// image rows [y0, y1) only (a tile-row band of a multi-GPU frame); `bgr` is the whole image
void launch_bgr8_rows(const float4 *fb_tiled, uint8_t *bgr, int W, int y0, int y1, int tile_w, int tile_h,
                      cudaStream_t st) {
  const int quads = (W / 4) * (y1 - y0);
  if (quads <= 0) return;
  ++g_launches;
  bgr8_kernel<<<ceil_div(quads, 256), 256, 0, st>>>(fb_tiled, bgr, W, y0, y1, tile_w, tile_h, W / tile_w);
}

}  // namespace gsplat
