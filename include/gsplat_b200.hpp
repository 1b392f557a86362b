// gsplat_b200.hpp — thin C++ shim over the C ABI (gsplat_b200.h) with the public surface of the
// reference's splat::IpuSplatter (include/splat/ipu_rasteriser.hpp:20-55), so that a host
// application written against IpuSplatter + GraphManager::execute changes one type name:
//
//   reference (src/main/splat.cpp:53-66,168,199,259-265)      here
//   ------------------------------------------------------   ---------------------------------------
//   IpuSplatter(gaussians, TiledFramebuffer&, noAMP)          B200Splatter(gaussians, n, width, height)
//   gm.compileOrLoad / gm.prepareEngine                       (constructor: device buffers + upload)
//   updateModelView(glm::mat4) / updateProjection(glm::mat4)  same names; any type with 16 contiguous
//   updateFocalLengths(fov, scaleDivisor)                     floats in glm order (glm::value_ptr)
//   gm.execute(*ipuSplatter)                                  execute()
//   getFrameBuffer(cv::Mat&)                                  getFrameBuffer(uint8_t* bgr, pitch)
//   getIPUHistogram(std::vector<uint32_t>&)                   same
//
// Errors: the reference throws std::runtime_error / std::logic_error from ipu_utils
// (include/ipu/ipu_utils.hpp:113-116,218-220,335-337); the shim throws std::runtime_error carrying
// gsplat_last_error().  Header-only; link with -lgsplat_b200.  No glm / OpenCV dependency.
#pragma once

#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "gsplat_b200.h"

namespace splat {

// The reference's Gaussian3D record (include/splat/ipu_geometry.hpp:305-311): 16 floats.
struct Gaussian3DRecord {
  float mean[4];    // x, y, z, w(=1)
  float colour[4];  // r, g, b, raw opacity
  float rot[4];     // quaternion w, x, y, z (rot_0 = w)
  float scale[3];   // log-scales
  float gid;        // index + 1; <= 0 marks an empty slot
};
static_assert(sizeof(Gaussian3DRecord) == 64, "must match the reference's 16-float record");

class B200Splatter {
public:
  // `gaussians`: n records; geometry defaults are the reference's compile-time constants
  // (tile_config.hpp:5-15).  device = CUDA ordinal.
  B200Splatter(const Gaussian3DRecord *gaussians, std::size_t n, int width = 1280, int height = 720, int device = 0,
               bool collectTimings = false) {
    gsplat_config cfg;
    gsplat_config_default(&cfg);
    cfg.device = device;
    cfg.width = width;
    cfg.height = height;
    cfg.collect_timings = collectTimings ? 1 : 0;
    if (gsplat_create(&cfg, &ctx_) != GSPLAT_OK) throw std::runtime_error(gsplat_last_error(nullptr));
    width_ = width; height_ = height;
    numTiles_ = (width / cfg.tile_w) * (height / cfg.tile_h);
    setIdentity(modelView_); setIdentity(projection_);
    try {
      check(gsplat_upload_scene(ctx_, reinterpret_cast<const float *>(gaussians), n));
    } catch (...) {
      gsplat_destroy(ctx_);
      ctx_ = nullptr;
      throw;
    }
  }
  explicit B200Splatter(const std::vector<Gaussian3DRecord> &g, int width = 1280, int height = 720, int device = 0)
      : B200Splatter(g.data(), g.size(), width, height, device) {}
  ~B200Splatter() { if (ctx_) gsplat_destroy(ctx_); }
  B200Splatter(const B200Splatter &) = delete;
  B200Splatter &operator=(const B200Splatter &) = delete;

  // ipu_rasteriser.cpp:86-110.  M is anything whose address is 16 contiguous floats in glm
  // (column-major) order: glm::mat4, float[16], std::array<float,16>.
  template <class M> void updateModelView(const M &mv) { copy16(modelView_, mv); }
  template <class M> void updateProjection(const M &mp) { copy16(projection_, mp); }
  void updateFocalLengths(float fx, float fy) { fov_ = fx; scaleDiv_ = fy; }

  // One "project" program run (ipu_rasteriser.cpp:393-399,408-420).  Asynchronous.
  void execute() {
    check(gsplat_set_camera(ctx_, modelView_, projection_, fov_, scaleDiv_));
    check(gsplat_render(ctx_));
  }
  void waitForCompletion() { check(gsplat_sync(ctx_)); }

  // ipu_rasteriser.cpp:115-144: H rows of BGR8, `pitch` bytes apart (cv::Mat: data, step).
  void getFrameBuffer(std::uint8_t *bgr, std::size_t pitch) const {
    check(gsplat_read_framebuffer_bgr8(ctx_, bgr, pitch));
  }
  void getFrameBuffer(std::vector<std::uint8_t> &bgr) const {
    bgr.resize(std::size_t(width_) * height_ * 3);
    getFrameBuffer(bgr.data(), std::size_t(width_) * 3);
  }
  // tile-major float RGBA, the outputFramebuffer stream (ipu_rasteriser.cpp:280-285)
  void getTiledFrameBuffer(std::vector<float> &fb) const {
    fb.resize(std::size_t(width_) * height_ * 4);
    check(gsplat_read_framebuffer_tiled_f32(ctx_, fb.data()));
  }
  // ipu_rasteriser.cpp:104-106
  void getIPUHistogram(std::vector<std::uint32_t> &counts) const {
    counts.resize(numTiles_);
    check(gsplat_read_tile_counts(ctx_, counts.data()));
  }

  // Frame stream (three slots): the render loop's imagePtr / imagePtrBuffered swap
  // (src/main/splat.cpp:215-226,281-283) with the device-to-host copy inside the pipeline.
  // executeStreamed() returns a ticket at once; acquireFrame(ticket, ...) waits for that frame and
  // points into pinned host memory that stays valid until the third-next executeStreamed().
  std::uint64_t executeStreamed() {
    check(gsplat_set_camera(ctx_, modelView_, projection_, fov_, scaleDiv_));
    std::uint64_t t = 0;
    check(gsplat_render_streamed(ctx_, &t));
    return t;
  }
  void acquireFrame(std::uint64_t ticket, const std::uint8_t **bgr, const std::uint32_t **counts = nullptr) {
    check(gsplat_acquire_frame(ctx_, ticket, bgr, counts));
  }

  // frames the context keeps in flight on the device (gsplat_config.concurrent_frames resolved)
  int concurrentFrames() const { return gsplat_concurrent_frames(ctx_); }

  // multi-GPU tile bands (no reference equivalent: single IPU, src/main/splat.cpp:41-46)
  void setBand(int rowBegin, int rowEnd) { check(gsplat_set_band(ctx_, rowBegin, rowEnd)); }
  gsplat_frame_stats stats() const {
    gsplat_frame_stats s;
    check(gsplat_get_frame_stats(ctx_, &s));
    return s;
  }
  gsplat_ctx *handle() const { return ctx_; }
  int width() const { return width_; }
  int height() const { return height_; }
  int numTiles() const { return numTiles_; }

private:
  void check(int rc) const {
    if (rc < 0) throw std::runtime_error(std::string("gsplat: ") + gsplat_last_error(ctx_));
  }
  static void setIdentity(float *m) {
    std::memset(m, 0, 16 * sizeof(float));
    m[0] = m[5] = m[10] = m[15] = 1.f;
  }
  template <class M> static void copy16(float *dst, const M &m) {
    static_assert(sizeof(M) == 16 * sizeof(float), "expected a 4x4 float matrix");
    std::memcpy(dst, &m, 16 * sizeof(float));
  }
  gsplat_ctx *ctx_ = nullptr;
  int width_ = 0, height_ = 0, numTiles_ = 0;
  float modelView_[16], projection_[16];
  float fov_ = 0.6981317f /* radians(40), src/main/splat.cpp:170 */, scaleDiv_ = 0.1f /* lambda1/10, :262 */;
};

}  // namespace splat
