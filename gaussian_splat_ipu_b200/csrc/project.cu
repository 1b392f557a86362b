// project.cu — K1: MVP transform, EWA 3D->2D covariance, bbox radius, render predicate and
// converged tile rectangle for every Gaussian.
//
// This translation unit is compiled with -fmad=false: every product and sum below is a
// separately rounded IEEE f32 operation in the order glm 0.9.9.9 evaluates it on the host
// (the reference's codelet is plain glm arithmetic, include/splat/ipu_geometry.hpp:315-383), so
// K1 is bit-identical to the host-compiled reference, exp() included (glibc's expf algorithm is
// restated in gsplat_internal.h).  The camera-independent part (ComputeCov3D) is cached per scene and
// scale divisor (cov3_kernel); K1 itself reads 64 B and writes 60 B per drawn Gaussian.
#include "gsplat_internal.h"

namespace gsplat {

struct M3 { float c[3][3]; };   // glm column-major: c[col][row]

// glm type_mat3x3.inl:486-518: Result[c][k] = A[0][k]*B[c][0] + A[1][k]*B[c][1] + A[2][k]*B[c][2]
__device__ __forceinline__ M3 m3_mul(const M3 &a, const M3 &b) {
  M3 r;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int k = 0; k < 3; ++k)
      r.c[c][k] = a.c[0][k] * b.c[c][0] + a.c[1][k] * b.c[c][1] + a.c[2][k] * b.c[c][2];
  return r;
}
__device__ __forceinline__ M3 m3_t(const M3 &a) {
  M3 r;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int k = 0; k < 3; ++k) r.c[c][k] = a.c[k][c];
  return r;
}
__device__ __forceinline__ float min_ref(float a, float b) { return a < b ? a : b; }   // ipu_geometry.hpp:329-331
__device__ __forceinline__ float max_ref(float a, float b) { return a > b ? a : b; }   // ipu_geometry.hpp:325-327
template <int THREADS>
__device__ __forceinline__ uint32_t block_sum(uint32_t v, uint32_t *smem /* THREADS/32 */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  uint32_t t = 0;
  if (warp == 0) {
    t = lane < THREADS / 32 ? smem[lane] : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  }
  return t;   // valid in warp 0
}

// Clamp a tile rectangle to the grid and to this context's row band; returns tiles touched.
__device__ __forceinline__ uint32_t clamp_rect(int &c0, int &c1, int &r0, int &r1, const FrameParams &fp) {
  c0 = c0 < 0 ? 0 : c0;
  c1 = c1 > fp.tiles_across - 1 ? fp.tiles_across - 1 : c1;
  r0 = r0 < 0 ? 0 : r0;
  r1 = r1 > fp.tiles_down - 1 ? fp.tiles_down - 1 : r1;
  int b0 = r0 < fp.row_begin ? fp.row_begin : r0;
  int b1 = r1 > fp.row_end - 1 ? fp.row_end - 1 : r1;
  if (c1 < c0 || b1 < b0) return 0;
  return (uint32_t)(c1 - c0 + 1) * (uint32_t)(b1 - b0 + 1);
}

__device__ __forceinline__ void
project_body(const SceneSoA &s, const FrameParams &fp, const Projected &out, uint32_t *__restrict__ drawn_counter,
             float *__restrict__ dbg_rec8, int32_t *__restrict__ dbg_rect4) {
  // the frame's tile-row band, for the kernels behind K1 in the (captured) frame: only K1's
  // argument block is refreshed per frame, so a band change needs no re-capture (drawn_counter is
  // word SC_V of the frame scalars; the band words sit at BAND_WORD_OFFSET from it)
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    drawn_counter[BAND_WORD_OFFSET] = (uint32_t)fp.row_begin;
    drawn_counter[BAND_WORD_OFFSET + 1] = (uint32_t)fp.row_end;
  }
  __shared__ uint32_t red[K1_THREADS / 32];
  const uint32_t i = blockIdx.x * K1_THREADS + threadIdx.x;
  uint32_t touched = 0;
  if (i < s.n) {
    const float4 p = __ldg(&s.pos_c[i]);
    const bool live = (p.x == p.x);   // empty slots (gid <= 0, codelets.cpp:456) carry a NaN position
    int c0 = 0, c1 = -1, r0 = 0, r1 = -1;
    uint2 rect_out = make_uint2(RECT_EMPTY, RECT_EMPTY);
    float px = 0.f, py = 0.f, cz = 0.f, cova = 0.f, covb = 0.f, covc = 0.f, rad = 0.f;
    bool drawn = false;
    if (live) {
      const float4 ca = __ldg(&s.cov_a[i]), cb = __ldg(&s.cov_b[i]);
      const float *m = fp.mvp;
      // clip = mvp * (x,y,z,1): glm type_mat4x4.inl:561-572, (m0*v0 + m1*v1) + (m2*v2 + m3*v3)
      float clip[4];
#pragma unroll
      for (int k = 0; k < 4; ++k)
        clip[k] = (m[0 + k] * p.x + m[4 + k] * p.y) + (m[8 + k] * p.z + m[12 + k] * 1.0f);
      // Viewport::clipSpaceToViewport, include/splat/viewport.hpp:21-35
      const float sdiv = .5f / clip[3];
      px = clip[0] * sdiv; py = clip[1] * sdiv;
      px += .5f; py += .5f;
      px *= fp.img_w; py *= fp.img_h;
      px += 0.0f; py += 0.0f;
      cz = clip[2];
      // ---- Gaussian3D::ComputeCov3D (ipu_geometry.hpp:315-323): camera independent, cached by
      // cov3_kernel below for the frame's scale divisor
      M3 cov3;
      cov3.c[0][0] = ca.x; cov3.c[0][1] = ca.y; cov3.c[0][2] = ca.z;
      cov3.c[1][0] = ca.w; cov3.c[1][1] = cb.x; cov3.c[1][2] = cb.y;
      cov3.c[2][0] = cb.z; cov3.c[2][1] = cb.w; cov3.c[2][2] = p.w;

      // ---- Gaussian3D::ComputeCov2D, ipu_geometry.hpp:333-383.  mv is the FULL projection*view
      // product and tan_fovy == tan_fovx (reference quirks, kept): t == clip.
      float tx = clip[0], ty = clip[1];
      const float tz = clip[2];
      const float lim = 1.3f * fp.tanfov;
      const float txtz = tx / tz, tytz = ty / tz;
      tx = min_ref(lim, max_ref(-lim, txtz)) * tz;
      ty = min_ref(lim, max_ref(-lim, tytz)) * tz;
      M3 J;
      J.c[0][0] = fp.focal_x / tz; J.c[0][1] = 0.0f; J.c[0][2] = -(fp.focal_x * tx) / (tz * tz);
      J.c[1][0] = 0.0f; J.c[1][1] = fp.focal_y / tz; J.c[1][2] = -(fp.focal_y * ty) / (tz * tz);
      J.c[2][0] = 0.f; J.c[2][1] = 0.f; J.c[2][2] = 0.f;
      M3 Wm;
#pragma unroll
      for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int k = 0; k < 3; ++k) Wm.c[c][k] = m[c * 4 + k];
      const M3 T = m3_mul(Wm, J);
      M3 cov = m3_mul(m3_mul(m3_t(T), m3_t(cov3)), T);
      cov.c[0][0] += 0.3f;
      cov.c[1][1] += 0.3f;
      cova = cov.c[0][0]; covb = cov.c[0][1]; covc = cov.c[1][1];

      // ---- Gaussian2D::ComputeEigenvalues + GetBoundingBox, ipu_geometry.hpp:247-276
      const float det = cova * covc - covb * covb;
      const float mid = .5f * (cova + covc);
      const float lambda1 = mid + sqrtf(max_ref(0.1f, mid * mid - det));
      const float lambda2 = mid - sqrtf(max_ref(0.1f, mid * mid - det));
      const float radius = ceilf(3.f * sqrtf(max_ref(lambda1, lambda2)));
      const float bminx = px - radius, bminy = py - radius, bmaxx = px + radius, bmaxy = py + radius;
      const float ddx = bmaxx - bminx, ddy = bmaxy - bminy;
      // codelets.cpp:470 `length(diagonal) < guard`: sqrtf(x) < guard  <=>  x < guard_sq (set_camera), also for
      // x = +inf (false both ways) and NaN (false both ways); x is a sum of squares, never negative
      const bool within = (ddx * ddx + ddy * ddy) < fp.guard_sq;
      rad = (bmaxx - bminx) * 0.5f;
      // a6: guard band, clip z < 0 (codelets.cpp:493), mean anchored by some tile
      // (Bounds2f::contains over the union of tiles, ipu_geometry.hpp:163-165)
      const bool anchored = ceilf(px) >= 0.f && floorf(px) < fp.img_w && ceilf(py) >= 0.f && floorf(py) < fp.img_h;
      drawn = within && (cz < 0.0f) && anchored;
      if (drawn) {
        // a7: converged lattice rectangle from Bounds2f::clip, ipu_geometry.hpp:133-155
        // x / 32 and x * (1 / 32) are the same float (scaling by a power of two is exact)
        if (fp.inv_tile_w != 0.0f) { c0 = (int)floorf(floorf(bminx) * fp.inv_tile_w); c1 = (int)floorf(ceilf(bmaxx) * fp.inv_tile_w); }
        else { c0 = (int)floorf(floorf(bminx) / fp.tile_w); c1 = (int)floorf(ceilf(bmaxx) / fp.tile_w); }
        r0 = (int)floorf(floorf(bminy) / fp.tile_h); r1 = (int)floorf(ceilf(bmaxy) / fp.tile_h);
        int b0 = r0, b1 = r1;
        touched = clamp_rect(c0, c1, b0, b1, fp);
        // the rectangle reported for parity is the grid-clamped one; the band clamp only
        // restricts which rows this context emits
        r0 = b0; r1 = b1;
        int e0 = r0 < fp.row_begin ? fp.row_begin : r0, e1 = r1 > fp.row_end - 1 ? fp.row_end - 1 : r1;
        if (touched) {
          const float4 col = __ldg(&s.rgba[i]);
          // Gaussian2D::ComputeConicOpacity, ipu_geometry.hpp:278-286 (per Gaussian, hoisted out
          // of the pixel loop; det == 0 -> zero conic and zero opacity -> never blended)
          float kx = 0.f, ky = 0.f, kz = 0.f, kw = 0.f;
          if (det != 0.0f) {
            const float det_inv = 1.f / det;
            kx = covc * det_inv; ky = -covb * det_inv; kz = cova * det_inv; kw = col.w;
          }
          // K5 record (gsplat_internal.h): mean, conic and opacity exactly as the reference's blend
          // loop recomputes them per pixel (ComputeConicOpacity), plus two CONSERVATIVE culling aids:
          // the half extents of the axis-aligned box around the ellipse alpha >= 1/255, which K5's
          // warps test against their 8x4 pixel blocks, and qmax, the largest value of the quadratic
          // form d^T Q d at which alpha >= 1/255 is possible.  Pairs skipped through them are pairs
          // the reference loop `continue`s over (codelets.cpp:400-403).
          float qmax = __int_as_float(0xff800000);                     // -inf: no pixel ever passes
          float hx = __int_as_float(0xff800000), hy = hx;              // -inf: touches no block
          if (kw * 255.0f > 1.0f) {                  // otherwise alpha = op * exp(<= 0) < 1/255 everywhere
            const float L = __logf(kw * 255.0f);     // power >= -L  <=>  op * exp(power) >= 1/255
            // margin: K5 evaluates the form fused, the reference's power unfused, and for thin splats the
            // three terms cancel (each up to ~1e5 where the form itself is ~20)
            qmax = 2.0f * L * 1.01f + 0.5f;
            hx = hy = __int_as_float(0x7f800000);    // indefinite form: no extent, every block is visited
            if (det > 0.f && cova > 0.f && covc > 0.f) {
              // ellipse d^T Q d <= 2L with Q = cov^-1: half-width sqrt(2L cov.xx), half-height
              // sqrt(2L cov.yy) — from the covariance itself, NOT from the conic: C / (A C - B^2)
              // cancels catastrophically for thin splats and undershoots the extent
              hx = sqrtf(2.0f * L * cova) * 1.01f + 0.5f;
              hy = sqrtf(2.0f * L * covc) * 1.01f + 0.5f;
              if (!(hx == hx) || !(hy == hy)) hx = hy = __int_as_float(0x7f800000);
            }
          }
          float4 *r = out.rec + (size_t)i * 3;
          r[0] = make_float4(px, py, hx, hy);
          r[1] = make_float4(kx, ky, kz, qmax);
          r[2] = make_float4(kw, col.x, col.y, col.z);
          rect_out = make_uint2((uint32_t)c0 | ((uint32_t)c1 << 16), (uint32_t)e0 | ((uint32_t)e1 << 16));
        }
      }
    }
    // depth key: drawn Gaussians have clip z < 0 (codelets.cpp:493); ascending z (:314) ==
    // descending magnitude == ascending ~bits(z); everything else sorts last
    out.depth_key[i] = touched ? ~__float_as_uint(cz) : DEPTH_KEY_CULLED;
    out.rect[i] = rect_out;   // RECT_EMPTY (c0 > c1: covers no tile) unless the Gaussian is drawn
    if (dbg_rec8) {
      float4 *d = reinterpret_cast<float4 *>(dbg_rec8 + (size_t)i * 8);
      d[0] = make_float4(px, py, cz, cova);
      d[1] = make_float4(covb, covc, rad, drawn ? 1.f : 0.f);
      reinterpret_cast<int4 *>(dbg_rect4)[i] = make_int4(c0, c1, r0, r1);
    }
  }
  const uint32_t nd = block_sum<K1_THREADS>(touched ? 1u : 0u, red);
  if (threadIdx.x == 0 && nd) atomicAdd(drawn_counter, nd);
}

// Two entry points of K1's body.  project_kernel: 40 registers, 6 CTAs per SM.  project_kernel_dense:
// capped at 32 registers (a few spilled words), 8 CTAs = all 64 warps per SM: K1 alone 0.150 -> 0.143 ms
// (0.85 of HBM), but it then leaves no room for another lane's kernels beside it — with several
// frames in flight 1504 against 1604 frames/s (profiles/ab_k5_r2e.txt).  So a context that renders one
// frame at a time launches the dense form, the lanes of a dispatcher the other.
__global__ void __launch_bounds__(K1_THREADS)
project_kernel(SceneSoA s, FrameParams fp, Projected out, uint32_t *__restrict__ drawn_counter,
               float *__restrict__ dbg_rec8, int32_t *__restrict__ dbg_rect4) {
  project_body(s, fp, out, drawn_counter, dbg_rec8, dbg_rect4);
}
__global__ void __launch_bounds__(K1_THREADS, 8)
project_kernel_dense(SceneSoA s, FrameParams fp, Projected out, uint32_t *__restrict__ drawn_counter,
                     float *__restrict__ dbg_rec8, int32_t *__restrict__ dbg_rect4) {
  project_body(s, fp, out, drawn_counter, dbg_rec8, dbg_rect4);
}

// Parity entry (gsplat_bin_projected): take depth / rect / drawn from the caller instead of K1.
__global__ void __launch_bounds__(K1_THREADS)
inject_kernel(const float *__restrict__ depth, const int32_t *__restrict__ rect4,
              const uint8_t *__restrict__ drawn, uint32_t n, FrameParams fp, Projected out,
              uint32_t *__restrict__ drawn_counter) {
  __shared__ uint32_t red[K1_THREADS / 32];
  if (blockIdx.x == 0 && threadIdx.x == 0) {   // as K1: publish the band
    drawn_counter[BAND_WORD_OFFSET] = (uint32_t)fp.row_begin;
    drawn_counter[BAND_WORD_OFFSET + 1] = (uint32_t)fp.row_end;
  }
  const uint32_t i = blockIdx.x * K1_THREADS + threadIdx.x;
  uint32_t touched = 0;
  if (i < n) {
    float z = 0.f;
    if (drawn[i]) {
      const int4 r = reinterpret_cast<const int4 *>(rect4)[i];
      int c0 = r.x, c1 = r.y, r0 = r.z, r1 = r.w;
      touched = clamp_rect(c0, c1, r0, r1, fp);
      const int e0 = r0 < fp.row_begin ? fp.row_begin : r0, e1 = r1 > fp.row_end - 1 ? fp.row_end - 1 : r1;
      if (touched) {
        z = depth[i];
        out.rect[i] = make_uint2((uint32_t)c0 | ((uint32_t)c1 << 16), (uint32_t)e0 | ((uint32_t)e1 << 16));
      }
    }
    out.depth_key[i] = touched ? ~__float_as_uint(z) : DEPTH_KEY_CULLED;
    if (!touched) out.rect[i] = make_uint2(RECT_EMPTY, RECT_EMPTY);
  }
  const uint32_t nd = block_sum<K1_THREADS>(touched ? 1u : 0u, red);
  if (threadIdx.x == 0 && nd) atomicAdd(drawn_counter, nd);
}

// Scene upload: reference AoS records (16 floats) -> SoA (56 live bytes; the covariance cache is
// filled by cov3_kernel).
__global__ void aos_to_soa_kernel(const float4 *__restrict__ aos, SceneSoA s) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= s.n) return;
  const float4 mean = aos[(size_t)i * 4 + 0], col = aos[(size_t)i * 4 + 1];
  const float4 rot = aos[(size_t)i * 4 + 2], sg = aos[(size_t)i * 4 + 3];
  const bool live = sg.w > 0.f;   // gid > 0
  s.pos_c[i] = make_float4(live ? mean.x : __int_as_float(0x7fc00000), mean.y, mean.z, 0.f);
  s.quat[i] = rot;
  s.rgba[i] = col;
  s.scl[i] = make_float4(sg.x, sg.y, sg.z, 0.f);
}

// Gaussian3D::ComputeCov3D for every Gaussian (ipu_geometry.hpp:315-323, after codelets.cpp:463's
// g.scale / fxy[1]): cov3 = R S S^T R^T in glm's evaluation order, exp() as the host libm's.  Runs
// when the scene or the scale divisor changes, not per frame.
__global__ void __launch_bounds__(K1_THREADS)
cov3_kernel(SceneSoA s, float scale_div) {
  // the exp table is indexed per thread: from constant memory every distinct index of a warp is a
  // separate replay, from shared memory it is one (at worst two-way conflicted) access
  __shared__ unsigned long long exp_tab[32];
  if (threadIdx.x < 32) exp_tab[threadIdx.x] = EXPF_TAB[threadIdx.x];
  __syncthreads();
  const uint32_t i = blockIdx.x * K1_THREADS + threadIdx.x;
  if (i >= s.n) return;
  const float4 sl = s.scl[i];
  const float4 q = s.quat[i];
  // g.scale / fxy[1] (codelets.cpp:463); x / 1.0f is x exactly
  float sc0 = sl.x, sc1 = sl.y, sc2 = sl.z;
  if (scale_div != 1.0f) { sc0 = sc0 / scale_div; sc1 = sc1 / scale_div; sc2 = sc2 / scale_div; }
  // glm normalize + mat3_cast
  float w = q.x, x = q.y, y = q.z, z = q.w;
  const float len = sqrtf((w * w + x * x) + (y * y + z * z));
  if (len <= 0.0f) { w = 1.f; x = 0.f; y = 0.f; z = 0.f; }
  else { const float inv = 1.0f / len; w = w * inv; x = x * inv; y = y * inv; z = z * inv; }
  const float qxx = x * x, qyy = y * y, qzz = z * z, qxz = x * z, qxy = x * y, qyz = y * z;
  const float qwx = w * x, qwy = w * y, qwz = w * z;
  M3 R;
  R.c[0][0] = 1.f - 2.f * (qyy + qzz); R.c[0][1] = 2.f * (qxy + qwz); R.c[0][2] = 2.f * (qxz - qwy);
  R.c[1][0] = 2.f * (qxy - qwz); R.c[1][1] = 1.f - 2.f * (qxx + qzz); R.c[1][2] = 2.f * (qyz + qwx);
  R.c[2][0] = 2.f * (qxz + qwy); R.c[2][1] = 2.f * (qyz - qwx); R.c[2][2] = 1.f - 2.f * (qxx + qyy);
  M3 S;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int k = 0; k < 3; ++k) S.c[c][k] = 0.f;
  S.c[0][0] = exp_ref(sc0, exp_tab); S.c[1][1] = exp_ref(sc1, exp_tab); S.c[2][2] = exp_ref(sc2, exp_tab);
  const M3 cov3 = m3_mul(m3_mul(m3_mul(R, S), m3_t(S)), m3_t(R));
  s.cov_a[i] = make_float4(cov3.c[0][0], cov3.c[0][1], cov3.c[0][2], cov3.c[1][0]);
  s.cov_b[i] = make_float4(cov3.c[1][1], cov3.c[1][2], cov3.c[2][0], cov3.c[2][1]);
  s.pos_c[i].w = cov3.c[2][2];
}

void launch_cov3(const SceneSoA &scene, float scale_div, cudaStream_t st) {
  if (!scene.n) return;
  ++g_launches;
  cov3_kernel<<<ceil_div(scene.n, K1_THREADS), K1_THREADS, 0, st>>>(scene, scale_div);
}

void launch_aos_to_soa(const float *aos16, SceneSoA scene, cudaStream_t st) {
  if (!scene.n) return;
  ++g_launches;
  aos_to_soa_kernel<<<ceil_div(scene.n, 256), 256, 0, st>>>(reinterpret_cast<const float4 *>(aos16), scene);
}

bool is_project_kernel(const void *func) {
  return func == reinterpret_cast<const void *>(&project_kernel) || func == reinterpret_cast<const void *>(&project_kernel_dense);
}

void launch_project(const SceneSoA &scene, const FrameParams &fp, Projected out, uint32_t *drawn_counter,
                    float *dbg_rec8, int32_t *dbg_rect4, bool dense, cudaStream_t st) {
  ++g_launches;   // an empty scene still launches one CTA: K1 publishes the frame's band
  const int grid = scene.n ? ceil_div(scene.n, K1_THREADS) : 1;
  if (dense) project_kernel_dense<<<grid, K1_THREADS, 0, st>>>(scene, fp, out, drawn_counter, dbg_rec8, dbg_rect4);
  else project_kernel<<<grid, K1_THREADS, 0, st>>>(scene, fp, out, drawn_counter, dbg_rec8, dbg_rect4);
}

void launch_inject(const float *depth, const int32_t *rect4, const uint8_t *drawn, uint32_t n,
                   const FrameParams &fp, Projected out, uint32_t *drawn_counter, cudaStream_t st) {
  ++g_launches;
  inject_kernel<<<n ? ceil_div(n, K1_THREADS) : 1, K1_THREADS, 0, st>>>(depth, rect4, drawn, n, fp, out, drawn_counter);
}

}  // namespace gsplat
