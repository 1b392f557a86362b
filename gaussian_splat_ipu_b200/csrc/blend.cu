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
// Mapping: one CTA per tile, one warp per 32x4 pixel strip, 4 pixels (one column of the strip)
// per thread held in registers; the list is staged through shared memory in batches of
// blockDim.x records gathered by index (idx 4 B + record 36 B per duplicate, once per tile).
// Stores to the tile-major framebuffer are 512 B contiguous per warp row.
#include "gsplat_internal.h"

namespace gsplat {

constexpr int PX = 4;            // pixels per thread (rows of a strip)
constexpr int BLEND_MAX_WARPS = 8;

__global__ void __launch_bounds__(BLEND_MAX_WARPS * 32)
blend_kernel(Projected p, const uint32_t *__restrict__ vals, const uint2 *__restrict__ ranges,
             int tiles_across, int row_begin, int tile_h, float4 *__restrict__ fb,
             uint32_t *__restrict__ counts) {
  __shared__ float4 sA[BLEND_MAX_WARPS * 32];   // mean.x, mean.y, conic.x, conic.y
  __shared__ float4 sB[BLEND_MAX_WARPS * 32];   // conic.z, opacity, r, g
  __shared__ float sC[BLEND_MAX_WARPS * 32];    // b
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t nthreads = blockDim.x;
  const uint32_t tile = blockIdx.x + (uint32_t)row_begin * tiles_across;
  const uint2 range = ranges[tile];
  const uint32_t n = range.y - range.x;
  if (tid == 0) counts[tile] = n;

  const uint32_t tcol = tile % tiles_across, trow = tile / tiles_across;
  const float xf = (float)(tcol * 32 + lane);
  const float y0 = (float)(trow * tile_h + warp * PX);

  float T[PX], Cr[PX], Cg[PX], Cb[PX], Ca[PX];
#pragma unroll
  for (int k = 0; k < PX; ++k) { T[k] = 1.f; Cr[k] = Cg[k] = Cb[k] = Ca[k] = 0.f; }
  uint32_t live = (1u << PX) - 1;   // bit k: pixel k has not hit the transmittance stop

  for (uint32_t base = 0; base < n; base += nthreads) {
    const uint32_t j = base + tid;
    if (j < n) {
      const uint32_t g = vals[range.x + j];
      sA[tid] = __ldg(&p.a[g]);
      sB[tid] = __ldg(&p.b[g]);
      sC[tid] = __ldg(&p.c[g].x);
    }
    __syncthreads();
    const uint32_t m = min(nthreads, n - base);
    if (live) {
      for (uint32_t jj = 0; jj < m; ++jj) {
        const float4 a = sA[jj];
        const float4 b = sB[jj];
        const float dx = a.x - xf;
        const float hx = -0.5f * a.z * dx * dx;   // -0.5 A dx^2
        const float bx = -a.w * dx;               // -B dx
        const float hc = -0.5f * b.x;             // -0.5 C
        bool any = false;
        float alpha[PX];
#pragma unroll
        for (int k = 0; k < PX; ++k) {
          const float dy = a.y - (y0 + (float)k);
          const float power = hx + dy * (bx + hc * dy);
          float al = fminf(0.99f, b.y * __expf(power));
          if (power > 0.f || al < (1.0f / 255.0f) || !((live >> k) & 1u)) al = 0.f;
          alpha[k] = al;
          any |= al > 0.f;
        }
        if (any) {
          const float cb = sC[jj];
#pragma unroll
          for (int k = 0; k < PX; ++k) {
            if (alpha[k] > 0.f) {
              const float test_T = T[k] * (1.f - alpha[k]);
              if (test_T < 0.0001f) {
                live &= ~(1u << k);
              } else {
                const float w = alpha[k] * T[k];
                Cr[k] += b.z * w; Cg[k] += b.w * w; Cb[k] += cb * w; Ca[k] += b.y * w;
                T[k] = test_T;
              }
            }
          }
          if (!live) break;
        }
      }
    }
    // all pixels of the tile stopped -> the rest of the list cannot contribute
    if (__syncthreads_and(live == 0)) break;
  }

  float4 *dst = fb + (size_t)tile * (32 * tile_h) + (warp * PX) * 32 + lane;
#pragma unroll
  for (int k = 0; k < PX; ++k) dst[k * 32] = make_float4(Cr[k], Cg[k], Cb[k], Ca[k]);
}

void launch_blend(const Projected &p, const uint32_t *sorted_vals, const uint2 *ranges, const FrameParams &fp,
                  float4 *fb_tiled, uint32_t *counts, cudaStream_t st) {
  const int rows = fp.row_end - fp.row_begin;
  if (rows <= 0) return;
  const int tile_h = (int)fp.tile_h;
  blend_kernel<<<rows * fp.tiles_across, (tile_h / PX) * 32, 0, st>>>(p, sorted_vals, ranges, fp.tiles_across,
                                                                     fp.row_begin, tile_h, fb_tiled, counts);
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
  bgr8_kernel<<<ceil_div(quads, 256), 256, 0, st>>>(fb_tiled, bgr, W, H, tile_w, tile_h, W / tile_w);
}

}  // namespace gsplat
