// ref_host_driver.cpp — TEST INFRASTRUCTURE (see pyoracle.py).  Thin C driver over the reference's OWN
// host-side input code, compiled unchanged from where it lies under /root/reference by
// oracle/build_ref.sh into oracle/_ref/libhost_ref.so:
//   src/splat/file_io.cpp   loadPoints / loadPlyFile / fillPlyProperties (happly, include/happly.h)
//   src/splat/camera.cpp    lookAtBoundingBox
//   src/splat/geometry.cpp  fitFrustumToBoundingBox
// plus glm for the orbit / UI view.  It pins gaussian_splat_ipu_b200/scene.py and apps/splat_b200.cpp
// (PLY loading, Gaussian3D conversion, cameras) to the reference, bit for bit.
// The only lines restated here are the ones that live in the reference's main(), which needs
// Poplar / OpenCV / Boost and cannot be compiled: recentre + z-negate (src/main/splat.cpp:93-100),
// the PLY -> Gaussian3D loop (:136-163), the bbox-to-eye-space step of the projection (:186-195),
// mvpStart (:235-241) and the UI view (:312-314) — each calls the reference's functions.
#include <splat/file_io.hpp>
#include <splat/camera.hpp>
#include <splat/geometry.hpp>
#include <glm/gtc/matrix_transform.hpp>
#include <glm/gtc/type_ptr.hpp>

#include <cstring>
#include <string>
#include <vector>

namespace {
struct HostScene {
  splat::Ply ply;
  splat::Points pts;
  splat::Bounds3f bb;           // after recentre + z-negate
  std::string error;
};
void put(const glm::mat4 &m, float *out16) { std::memcpy(out16, glm::value_ptr(m), 16 * sizeof(float)); }
}  // namespace

extern "C" {

// loadPoints (file_io.cpp:44-55) + Bounds3f + recentre / z-negate (splat.cpp:86-100)
void *refhost_open(const char *path) {
  auto *s = new HostScene;
  try {
    s->pts = splat::loadPoints(path, s->ply);
    splat::Bounds3f bb(s->pts);
    const auto bbCentre = bb.centroid();
    for (auto &v : s->pts) {
      v.p -= bbCentre;
      v.p.z = -v.p.z;
    }
    s->bb = splat::Bounds3f(s->pts);
  } catch (const std::exception &e) {
    s->error = e.what();
  }
  return s;
}
const char *refhost_error(void *h) { return static_cast<HostScene *>(h)->error.c_str(); }
void refhost_close(void *h) { delete static_cast<HostScene *>(h); }
size_t refhost_count(void *h) { return static_cast<HostScene *>(h)->pts.size(); }

// the 14 float properties fillPlyProperties pulls (file_io.cpp:62-77), n x 14 in that order
void refhost_properties(void *h, float *out) {
  auto *s = static_cast<HostScene *>(h);
  const splat::Ply::Property *props[14] = {&s->ply.x, &s->ply.y, &s->ply.z, &s->ply.f_dc[0], &s->ply.f_dc[1], &s->ply.f_dc[2],
                                           &s->ply.opacity, &s->ply.scale[0], &s->ply.scale[1], &s->ply.scale[2],
                                           &s->ply.rot[0], &s->ply.rot[1], &s->ply.rot[2], &s->ply.rot[3]};
  const size_t n = s->pts.size();
  for (size_t i = 0; i < n; ++i)
    for (int k = 0; k < 14; ++k) out[i * 14 + k] = props[k]->values[i];
}

// the Gaussian3D records of src/main/splat.cpp:136-163 (mean4, colour4, rot4, scale3, gid), n x 16
void refhost_gaussians(void *h, float *out16) {
  auto *s = static_cast<HostScene *>(h);
  const auto &ply = s->ply;
  const float SH_C0 = 0.28209479177387814f;
  for (std::size_t i = 0; i < s->pts.size(); ++i) {
    auto pt = s->pts[i].p;
    float *g = out16 + i * 16;
    g[0] = pt.x; g[1] = pt.y; g[2] = pt.z; g[3] = 1.f;
    glm::vec3 colour = {SH_C0 * ply.f_dc[0].values[i], SH_C0 * ply.f_dc[1].values[i], SH_C0 * ply.f_dc[2].values[i]};
    colour += 0.5f;
    colour = glm::max(colour, glm::vec3(0.f));
    g[4] = colour.x; g[5] = colour.y; g[6] = colour.z; g[7] = ply.opacity.values[i];
    g[8] = ply.rot[0].values[i]; g[9] = ply.rot[1].values[i]; g[10] = ply.rot[2].values[i]; g[11] = ply.rot[3].values[i];
    g[12] = ply.scale[0].values[i]; g[13] = ply.scale[1].values[i]; g[14] = ply.scale[2].values[i];
    g[15] = static_cast<float>(i) + 1.0f;
  }
}
void refhost_bounds(void *h, float *min3, float *max3) {
  auto *s = static_cast<HostScene *>(h);
  std::memcpy(min3, &s->bb.min, 12); std::memcpy(max3, &s->bb.max, 12);
}

// lookAtBoundingBox (camera.cpp:10-15)
void refhost_look_at_bounding_box(const float *min3, const float *max3, const float *up3, float scale, float *out16) {
  splat::Bounds3f bb(glm::make_vec3(min3), glm::make_vec3(max3));
  put(splat::lookAtBoundingBox(bb, glm::make_vec3(up3), scale), out16);
}
// fitFrustumToBoundingBox (geometry.cpp:9-24)
void refhost_fit_frustum(const float *min3, const float *max3, float fov, float aspect, float *out16) {
  splat::Bounds3f bb(glm::make_vec3(min3), glm::make_vec3(max3));
  put(splat::fitFrustumToBoundingBox(bb, fov, aspect), out16);
}
// the app's projection, src/main/splat.cpp:186-195
void refhost_app_projection(const float *min3, const float *max3, float fov, float aspect, float *out16) {
  splat::Bounds3f bb(glm::make_vec3(min3), glm::make_vec3(max3));
  auto modelView = splat::lookAtBoundingBox(bb, glm::vec3(0.f, 1.f, 1.f), 1.f);
  splat::Bounds3f bbInCamera(modelView * glm::vec4(bb.min, 1.f), modelView * glm::vec4(bb.max, 1.f));
  put(splat::fitFrustumToBoundingBox(bbInCamera, fov, aspect), out16);
}
// mvpStart (splat.cpp:235-241) rotated about y: the config-2 orbit
void refhost_orbit_view(float degrees, float *out16) {
  auto mvpStart = glm::mat4(1.0f);
  mvpStart[0][0] = -1.0f;
  mvpStart[1][1] = -0.09709989f;
  mvpStart[1][2] = -0.99527466f;
  mvpStart[2][1] = -0.99527466f;
  mvpStart[2][2] = 0.09709989f;
  mvpStart[3][2] = -5.1539507f;
  put(degrees == 0.f ? mvpStart : glm::rotate(mvpStart, glm::radians(degrees), glm::vec3(0.f, 1.f, 0.f)), out16);
}
// the UI view, src/main/splat.cpp:312-314
void refhost_ui_view(const float *modelView16, float envDeg, float envDeg2, float X, float Y, float Z, float *out16) {
  const glm::mat4 modelView = glm::make_mat4(modelView16);
  auto dynamicView = modelView * glm::rotate(glm::mat4(1.f), glm::radians(envDeg), glm::vec3(1.f, 0.f, 0.f));
  dynamicView = glm::rotate(dynamicView, glm::radians(envDeg2), glm::vec3(0.f, 1.f, 0.f));
  dynamicView = glm::translate(dynamicView, glm::vec3(X / 50.f, Y / 50.f, -Z / 20.f + 20.f));
  put(dynamicView, out16);
}
// glm::rotate(m, angle, axis): what the synthetic orbits use
void refhost_rotate(const float *m16, float angle, const float *axis3, float *out16) {
  put(glm::rotate(glm::make_mat4(m16), angle, glm::make_vec3(axis3)), out16);
}

}  // extern "C"
