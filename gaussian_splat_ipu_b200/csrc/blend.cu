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
// of the 32x20 tiles its rectangle covers, so the CTA *is* the semantic tile.
//
// Mapping: one CTA per tile (or column strip of it), one warp per 8 x 4 pixel BLOCK, one pixel per
// thread in registers.  The list is staged through shared memory in batches of records gathered
// by index; the staging thread also derives, per Gaussian, the rows and 8-pixel column blocks of
// the part on which alpha >= 1/255 is possible at all (extent of the ellipse
// power >= ln(1/(255 opacity)), conservative).  Each warp ballots over those extents and only
// walks the Gaussians that can touch its block: skipped pairs are exactly pairs the reference loop
// `continue`s over, so the result is unchanged while small splats cost ~nothing on the blocks
// they miss.
// exp() is ex2.approx.ftz on a log2e-prescaled conic.  FP32-pipe bound; no tensor cores (each
// pair has its own quadratic form).  Stores are four 128 B row segments per warp.
#include "gsplat_internal.h"

namespace gsplat {

unsigned long long g_launches = 0;

constexpr int BLEND_MAX_ROWS = 32;
constexpr float LOG2E = 1.4426950408889634f;
constexpr float ALPHA_MIN = 1.0f / 255.0f;

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Shared-memory batch, blockDim.x entries of each (dynamic, sized by the launcher):
//   sA float4: mean.x, mean.y, -0.5*log2e*A, -log2e*B     sB float4: -0.5*log2e*C, opacity, r, g
//   sC float2: b, conservative log2 threshold (skip when power2 < it)
//   sRows u16: row_lo | row_hi << 5 | mask of the part's 8-pixel column blocks << 10 ; 1: touches nothing
#ifndef GSPLAT_BLEND_BATCH
#define GSPLAT_BLEND_BATCH 2
#endif
constexpr uint32_t BLEND_BATCH = GSPLAT_BLEND_BATCH;   // list entries staged per thread and batch
constexpr size_t BLEND_SMEM_PER_THREAD = BLEND_BATCH * (16 + 16 + 8 + 2);

//
// Depth slices (see duplicate.cu): the list of a tile may arrive in several consecutive pieces, one
// launch each.  Between launches a pixel's state lives in the framebuffer (C) and in `tstate`
// (T, or -1 once the pixel has stopped); a tile whose pixels have all stopped raises
// tile_done[tile] so that later slices neither emit duplicates for it nor launch work on it.
// FIRST: start from T = 1, C = 0.  LAST: no state is saved.
template <bool COUNT, bool FIRST, bool LAST>
__global__ void __launch_bounds__(BLEND_MAX_ROWS * 32)
blend_kernel(Projected p, const uint32_t *__restrict__ vals, const uint2 *__restrict__ ranges,
             int tiles_across, int row_begin, int full_tile_h, int part_w, float4 *__restrict__ fb,
             float *__restrict__ tstate, TileState ts, const uint32_t *__restrict__ order,
             unsigned long long *__restrict__ pair_counters) {
  extern __shared__ __align__(16) unsigned char blend_smem_raw[];
  const long long t_begin = COUNT ? clock64() : 0;
  const uint32_t tid = threadIdx.x, lane = tid & 31, row = tid >> 5;
  const uint32_t nthreads = blockDim.x;
  float4 *sA = reinterpret_cast<float4 *>(blend_smem_raw);
  float4 *sB = sA + BLEND_BATCH * nthreads;
  float2 *sC = reinterpret_cast<float2 *>(sB + BLEND_BATCH * nthreads);
  uint16_t *sRows = reinterpret_cast<uint16_t *>(sC + BLEND_BATCH * nthreads);
  // a CTA blends a "part" of a semantic tile: all its rows, `part_w` = 32 / nsub of its columns (the
  // whole tile, or a column strip so that a crowded tile's list is walked by several SMs at
  // once); parts of one tile are neighbours in the grid.  A warp owns an 8 x 4 pixel block: a splat
  // a few pixels across then occupies a third of the lanes of the warps it reaches, where a
  // 32 x 1 pixel row would leave nine lanes in ten idle.
  const uint32_t nsub = (uint32_t)(32 / part_w), sub = blockIdx.x % nsub;
  const uint32_t cbn = (uint32_t)part_w / 8u;            // 8-pixel column blocks per part
  const uint32_t cb = row % cbn, rb = row / cbn;         // `row` = warp index: its block
  const uint32_t x_local = sub * (uint32_t)part_w + cb * 8u + (lane & 7u), y_local = rb * 4u + (lane >> 3);
  const bool valid = y_local < (uint32_t)full_tile_h;    // tile heights that are no multiple of 4
  // tiles are taken longest list first (tile_order_kernel) so the long ones do not start last
  const uint32_t tile = order[blockIdx.x / nsub] + (uint32_t)row_begin * tiles_across;
  const uint2 range = ranges[tile];
  const uint32_t n = range.y - range.x;
  const uint32_t part = tile * nsub + sub;
  if (!FIRST && (n == 0 || ts.part_done[part])) return;   // nothing new for this part: its state stands
  // a10 layout: tile-major, row-major inside the tile
  const size_t px = (size_t)tile * (32 * full_tile_h) + (size_t)(valid ? y_local : 0u) * 32 + x_local;

  const uint32_t tcol = tile % tiles_across, trow = tile / tiles_across;
  const float xp0 = (float)(tcol * 32u + sub * (uint32_t)part_w);   // first pixel column of the part
  const float xf = (float)(tcol * 32u + x_local);
  const float ty0 = (float)(trow * full_tile_h);
  const float yf = ty0 + (float)y_local;
  const int tile_h = full_tile_h;

  float T = 1.f, Cr = 0.f, Cg = 0.f, Cb = 0.f, Ca = 0.f;
  bool done = !valid;
  if (!FIRST && valid) {
    const float4 c = fb[px];
    Cr = c.x; Cg = c.y; Cb = c.z; Ca = c.w;
    T = tstate[px];
    done = T < 0.f;
  }
  uint32_t stop_at = done ? 0u : n;   // COUNT: list entries this pixel visits (reference semantics)
  unsigned long long n_blend = 0;  // COUNT

  const uint32_t batch = BLEND_BATCH * nthreads;   // list entries staged per barrier pair
  for (uint32_t base = 0; base < n; base += batch) {
    // ---- stage one batch: gather the record, prescale, derive the row range ----------------
#pragma unroll
    for (uint32_t sb = 0; sb < BLEND_BATCH; ++sb) {
    const uint32_t slot = sb * nthreads + tid, j = base + slot;
    if (j < n) {
      const uint32_t g = vals[range.x + j];
      const float4 *r = p.rec + (size_t)g * 3;   // 48 contiguous bytes: two sectors of one line
      const float4 a = __ldg(r);
      const float4 b = __ldg(r + 1);
      const float cb = __ldg(&r[2].x);
      const float A = a.z, B = a.w, C = b.x, op = b.y;
      uint32_t lo = 1, hi = 0, cmask = 0xfu;
      float thr2 = __int_as_float(0x7f800000);   // +inf: never passes
      if (op * 255.0f > 1.0f) {                  // otherwise alpha = op*exp(<=0) < 1/255 everywhere
        const float L2 = __log2f(op * 255.0f);   // power2 >= -L2  <=>  alpha >= 1/255
        thr2 = -L2 - 0.02f;
        lo = 0; hi = (uint32_t)tile_h - 1;
        const float det = A * C - B * B;
        if (det > 0.f && A > 0.f) {
          // ellipse d^T Q d <= 2 ln(255 op): half-height sqrt(2L * A / det); padded
          const float hy = sqrtf(2.0f * 0.6931472f * L2 * A / det) * 1.001f + 0.02f;
          const float flo = ceilf(a.y - hy - ty0), fhi = floorf(a.y + hy - ty0);
          if (fhi < 0.f || flo > (float)(tile_h - 1)) { lo = 1; hi = 0; }
          else { lo = flo > 0.f ? (uint32_t)flo : 0u; hi = fhi < (float)(tile_h - 1) ? (uint32_t)fhi : (uint32_t)tile_h - 1; }
          // Tighten to the rows the ellipse reaches INSIDE this tile's 32 columns: on row y the
          // quadratic form is smallest at dx* = clamp(-B dy / A, [dx of last column, dx of first]);
          // rows whose minimum already exceeds the bound (with margin) hold no pixel with
          // alpha >= 1/255.  The set of passing rows is an interval (convexity): trim both ends.
          const float bound = 2.0f * 0.6931472f * (L2 + 0.02f) * 1.001f + 1e-4f;
          const float dxa = a.x - (xp0 + (float)(part_w - 1)), dxb = a.x - xp0, invA = 1.0f / A;
          auto reaches = [&](uint32_t rrow) {
            const float dy = a.y - (ty0 + (float)rrow);
            const float dx = fminf(fmaxf(-B * dy * invA, dxa), dxb);
            return A * dx * dx + 2.0f * B * dx * dy + C * dy * dy <= bound;
          };
          while (lo <= hi && !reaches(lo)) ++lo;
          while (hi > lo && !reaches(hi)) --hi;
          if (lo > hi) { lo = 1; hi = 0; }
          // the 8-pixel column blocks of the part the ellipse reaches: half-width sqrt(2L * C / det)
          if (lo <= hi && C > 0.f) {
            const float hx = sqrtf(2.0f * 0.6931472f * L2 * C / det) * 1.001f + 0.02f;
            const float fl = ceilf(a.x - hx - xp0), fh = floorf(a.x + hx - xp0);
            if (fh < 0.f || fl > (float)(part_w - 1)) { lo = 1; hi = 0; }
            else if (fl == fl && fh == fh) {
              const uint32_t b0 = fl > 0.f ? (uint32_t)fl >> 3 : 0u;
              const uint32_t b1 = fh < (float)(part_w - 1) ? (uint32_t)fh >> 3 : (uint32_t)(part_w - 1) >> 3;
              cmask = ((2u << b1) - 1u) & ~((1u << b0) - 1u);
            }
          }
        }
      }
      sA[slot] = make_float4(a.x, a.y, -0.5f * LOG2E * A, -LOG2E * B);
      sB[slot] = make_float4(-0.5f * LOG2E * C, op, b.z, b.w);
      sC[slot] = make_float2(cb, thr2);
      sRows[slot] = (uint16_t)(lo > hi ? 1u : lo | (hi << 5) | (cmask << 10));
    }
    }
    __syncthreads();
    const uint32_t m = min(batch, n - base);
    if (!__all_sync(0xffffffffu, done)) {
      for (uint32_t c0 = 0; c0 < m; c0 += 32) {
        const uint32_t jj = c0 + lane;
        uint32_t rr = jj < m ? sRows[jj] : 1u;   // 1 = lo 1, hi 0: empty
        uint32_t mask = __ballot_sync(0xffffffffu, (rr & 31u) <= rb * 4u + 3u && ((rr >> 5) & 31u) >= rb * 4u &&
                                                       ((rr >> (10u + cb)) & 1u));
        // two list entries per turn: their loads, quadratic forms and ex2 overlap; they are then
        // applied to the pixel strictly in list order
        while (mask) {
          const uint32_t k0 = c0 + __ffs(mask) - 1;
          mask &= mask - 1;
          const bool two = mask != 0;
          const uint32_t k1 = two ? c0 + __ffs(mask) - 1 : k0;
          mask &= mask - 1;
          const float4 a0 = sA[k0], a1 = sA[k1];
          const float4 b0 = sB[k0], b1 = sB[k1];
          const float2 e0 = sC[k0], e1 = sC[k1];
          const float dx0 = a0.x - xf, dy0 = a0.y - yf, dx1 = a1.x - xf, dy1 = a1.y - yf;
          // power * log2e = dx*(hA*dx + nB*dy) + hC*dy*dy
          const float p0 = fmaf(b0.x * dy0, dy0, dx0 * fmaf(a0.w, dy0, a0.z * dx0));
          const float p1 = fmaf(b1.x * dy1, dy1, dx1 * fmaf(a1.w, dy1, a1.z * dx1));
          const bool v0 = p0 <= 0.f && p0 >= e0.y, v1 = two && p1 <= 0.f && p1 >= e1.y;
          const float al0 = v0 ? fminf(0.99f, b0.y * ex2_approx(p0)) : 0.f;
          const float al1 = v1 ? fminf(0.99f, b1.y * ex2_approx(p1)) : 0.f;
          if (!done && al0 >= ALPHA_MIN) {
            const float test_T = T * (1.f - al0);
            if (test_T < 0.0001f) {
              done = true;
              if (COUNT) stop_at = base + k0 + 1;
            } else {
              const float w = al0 * T;
              Cr = fmaf(b0.z, w, Cr); Cg = fmaf(b0.w, w, Cg); Cb = fmaf(e0.x, w, Cb); Ca = fmaf(b0.y, w, Ca);
              T = test_T;
              if (COUNT) ++n_blend;
            }
          }
          if (!done && al1 >= ALPHA_MIN) {
            const float test_T = T * (1.f - al1);
            if (test_T < 0.0001f) {
              done = true;
              if (COUNT) stop_at = base + k1 + 1;
            } else {
              const float w = al1 * T;
              Cr = fmaf(b1.z, w, Cr); Cg = fmaf(b1.w, w, Cg); Cb = fmaf(e1.x, w, Cb); Ca = fmaf(b1.y, w, Ca);
              T = test_T;
              if (COUNT) ++n_blend;
            }
          }
        }
        if (__all_sync(0xffffffffu, done)) break;
      }
    }
    // all pixels of the tile stopped -> the rest of the list cannot contribute
    if (__syncthreads_and(done)) break;
  }
  if (!LAST) {
    if (valid) tstate[px] = done ? -1.f : T;
    if (__syncthreads_and(done) && tid == 0) {
      ts.part_done[part] = 1;
      if (atomicAdd(&ts.parts_done[tile], 1u) + 1u == nsub) ts.tile_done[tile] = 1;   // last part of the tile to stop
    }
  }

  if (COUNT) {
    unsigned long long n_visit = stop_at;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      n_visit += __shfl_xor_sync(0xffffffffu, n_visit, o);
      n_blend += __shfl_xor_sync(0xffffffffu, n_blend, o);
    }
    if (lane == 0) {
      atomicAdd(&pair_counters[0], n_visit); atomicAdd(&pair_counters[1], n_blend);
      atomicAdd(&pair_counters[2 + 2 * (size_t)tile + 1], n_visit);            // per-tile profile tap
      atomicAdd(&pair_counters[2 + 2 * (size_t)tile], (unsigned long long)(clock64() - t_begin) / (unsigned)full_tile_h);
    }
  }
  if (valid) fb[px] = make_float4(Cr, Cg, Cb, Ca);
}

// ---- launch order of the tiles: longest list first ---------------------------------------------
// CTAs are dispatched in blockIdx order; a tile with a long list that starts late leaves the other
// SMs idle at the end of the launch.  One CTA buckets the band's tiles by list length (power-of-two
// class + 2 mantissa bits, longest class first; tiles that already stopped count as empty) and
// writes the permutation K5 indexes with blockIdx.  Order inside a class is arbitrary: it only
// changes scheduling, never results.
constexpr int ORDER_BUCKETS = 33 * 4;
__device__ __forceinline__ uint32_t order_bucket(uint32_t len) {
  if (len == 0) return ORDER_BUCKETS - 1;
  const uint32_t e = 31u - (uint32_t)__clz(len);                       // floor(log2 len)
  const uint32_t m = e >= 2 ? (len >> (e - 2)) & 3u : (len << (2 - e)) & 3u;
  return (31u - e) * 4u + (3u - m);                                     // longer lists -> smaller bucket
}
__global__ void __launch_bounds__(1024)
tile_order_kernel(const uint2 *__restrict__ ranges, int tiles_across, int row_begin, int ntiles,
                  const uint8_t *__restrict__ tile_done, uint32_t *__restrict__ order) {
  __shared__ uint32_t cnt[ORDER_BUCKETS];
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
    order[atomicAdd(&cnt[order_bucket(len)], 1u)] = (uint32_t)i;
  }
}

static void launch_tile_order(const uint2 *ranges, const FrameParams &fp, const uint8_t *tile_done, bool first,
                              uint32_t *order, cudaStream_t st) {
  const int ntiles = (fp.row_end - fp.row_begin) * fp.tiles_across;
  if (ntiles <= 0) return;
  ++g_launches;
  tile_order_kernel<<<1, 1024, 0, st>>>(ranges, fp.tiles_across, fp.row_begin, ntiles, first ? nullptr : tile_done, order);
}

template <bool COUNT, bool FIRST, bool LAST>
static void launch_blend_t(const Projected &p, const uint32_t *sorted_vals, const uint2 *ranges, const FrameParams &fp,
                           float4 *fb_tiled, float *tstate, const TileState &ts, int split, const uint32_t *order,
                           unsigned long long *pair_counters, cudaStream_t st) {
  const int rows = fp.row_end - fp.row_begin;
  if (rows <= 0) return;
  const int full_h = (int)fp.tile_h;
  const int part_w = 32 / split;                                  // columns per CTA (split is 1, 2 or 4)
  const int threads = (part_w / 8) * ((full_h + 3) / 4) * 32;    // one warp per 8 x 4 pixel block
  // dynamic smem sized for this part so that several CTAs fit per SM
  const size_t smem = (size_t)threads * BLEND_SMEM_PER_THREAD;
  ++g_launches;
  blend_kernel<COUNT, FIRST, LAST><<<rows * fp.tiles_across * split, threads, smem, st>>>(
      p, sorted_vals, ranges, fp.tiles_across, fp.row_begin, full_h, part_w, fb_tiled, tstate, ts, order, pair_counters);
}

template <bool COUNT, bool FIRST, bool LAST> static cudaError_t blend_attr() {
  return cudaFuncSetAttribute(blend_kernel<COUNT, FIRST, LAST>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              (int)(BLEND_MAX_ROWS * 32 * BLEND_SMEM_PER_THREAD));
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
                  float4 *fb_tiled, float *tstate, const TileState &ts, int split, uint32_t *order, bool first,
                  bool last, unsigned long long *pair_counters, cudaStream_t st) {
  launch_tile_order(ranges, fp, ts.tile_done, first, order, st);
#define GSPLAT_BLEND(C, F, L) launch_blend_t<C, F, L>(p, sorted_vals, ranges, fp, fb_tiled, tstate, ts, split, order, pair_counters, st)
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
bgr8_kernel(const float4 *__restrict__ fb, uint8_t *__restrict__ bgr, int W, int H, int tile_w, int tile_h,
            int tiles_across) {
  const int quads_per_row = W / 4;
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= quads_per_row * H) return;
  const int y = q / quads_per_row, x = (q % quads_per_row) * 4;
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
  const int quads = (W / 4) * H;
  ++g_launches;
  bgr8_kernel<<<ceil_div(quads, 256), 256, 0, st>>>(fb_tiled, bgr, W, H, tile_w, tile_h, W / tile_w);
}

}  // namespace gsplat
