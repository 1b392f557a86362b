// splat_b200 — host driver for the B200 splat core, the headless counterpart of the reference's
// `splat` executable (src/main/splat.cpp): load a Gaussian-splat PLY, build the Gaussian3D records
// and the cameras exactly as the app does, render an orbit through the drop-in boundary
// (include/gsplat_b200.hpp, the IpuSplatter-shaped shim) with the frame stream, write the last
// frame as a PPM and print the frame rate.  No Boost / OpenCV / spdlog / Poplar: plain C++17.
//
//   splat_b200 --input scene.ply [--width 1280] [--height 720] [--frames 32] [--scale-div 0.1]
//              [--out frame.ppm] [--device 0] [--cache scene.gsc] [--cache-only]
//   splat_b200 --synthetic 1000000 ...          (SURVEY §8d generator instead of a PLY)
//   splat_b200 --input scene.ply --ui-port 5000 interactive: frames streamed to remote_render_ui over
//                                               packetcomms, camera / fov / scale from the UI (splat.cpp:245-322)
//
// What replaces what (reference file:line):
//   PLY reading                 happly via src/splat/file_io.cpp:30-42,62-77: the SAME happly calls when built
//                               with -DGSPLAT_WITH_HAPPLY (apps/Makefile does when include/happly.h of the
//                               reference tree is reachable: in the reference's own tree it always is), else
//                               the built-in reader below; both give bit-identical records (tests/test_host_inputs.py)
//   tracing                     pvti tracepoints "splatter" / mvp_transform_ipu / ui_update (splat.cpp:69,210,258)
//                               -> NVTX ranges of the same names
//   PLY -> Gaussian3D           src/main/splat.cpp:136-163 (SH_C0 colour, raw opacity / rot / log-scale)
//   recentre + z-negate         src/main/splat.cpp:93-100
//   projection                  src/main/splat.cpp:186-195, src/splat/camera.cpp:10-15, src/splat/geometry.cpp:9-24
//   initial view + orbit        src/main/splat.cpp:235-241 (mvpStart), glm::rotate about y (config 2)
//   per-frame update + execute  src/main/splat.cpp:259-265
//   image hand-over             src/main/splat.cpp:215-226,281-283 (imagePtr / imagePtrBuffered swap)
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "gsplat_b200.hpp"
#ifdef GSPLAT_WITH_UI
#include "stream_server.hpp"   // B200StreamServer: the reference's InterfaceServer over packetcomms
#endif
#ifdef GSPLAT_WITH_HAPPLY
#include <happly.h>   // the reference's vendored loader (include/happly.h), not part of this repository
#endif
#if __has_include(<nvtx3/nvToolsExt.h>)
#include <nvtx3/nvToolsExt.h>   // header-only (dlopens the injection library when a tool is attached)
struct TraceScope {   // pvti::Tracepoint's role (src/main/splat.cpp:210-258)
  explicit TraceScope(const char *name) { nvtxRangePushA(name); }
  ~TraceScope() { nvtxRangePop(); }
};
#else
struct TraceScope { explicit TraceScope(const char *) {} };
#endif

namespace {

struct Mat4 {   // glm memory order: m[col][row]
  float m[4][4];
};
Mat4 identity() {
  Mat4 r{};
  for (int i = 0; i < 4; ++i) r.m[i][i] = 1.f;
  return r;
}
struct Vec3 {
  float x, y, z;
};
Vec3 sub(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
float dot(Vec3 a, Vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
Vec3 cross(Vec3 a, Vec3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
Vec3 normalize(Vec3 v) {
  const float s = 1.f / std::sqrt(dot(v, v));
  return {v.x * s, v.y * s, v.z * s};
}

// glm::lookAtRH, ext/matrix_transform.inl:153-173
Mat4 look_at(Vec3 eye, Vec3 center, Vec3 up) {
  const Vec3 f = normalize(sub(center, eye)), s = normalize(cross(f, up)), u = cross(s, f);
  Mat4 r = identity();
  r.m[0][0] = s.x; r.m[1][0] = s.y; r.m[2][0] = s.z;
  r.m[0][1] = u.x; r.m[1][1] = u.y; r.m[2][1] = u.z;
  r.m[0][2] = -f.x; r.m[1][2] = -f.y; r.m[2][2] = -f.z;
  r.m[3][0] = -dot(s, eye); r.m[3][1] = -dot(u, eye); r.m[3][2] = dot(f, eye);
  return r;
}

// glm::frustumRH_NO, ext/matrix_clip_space.inl:165-176
Mat4 frustum(float l, float r, float b, float t, float n, float f) {
  Mat4 m{};
  m.m[0][0] = (2.f * n) / (r - l);
  m.m[1][1] = (2.f * n) / (t - b);
  m.m[2][0] = (r + l) / (r - l);
  m.m[2][1] = (t + b) / (t - b);
  m.m[2][2] = -(f + n) / (f - n);
  m.m[2][3] = -1.f;
  m.m[3][2] = -(2.f * f * n) / (f - n);
  return m;
}

// glm::rotate(m, angle, axis), ext/matrix_transform.inl:18-46
Mat4 rotate(const Mat4 &m, float angle, Vec3 axis) {
  const float c = std::cos(angle), s = std::sin(angle);
  const Vec3 a = normalize(axis), t = {(1.f - c) * a.x, (1.f - c) * a.y, (1.f - c) * a.z};
  float r[3][3];
  r[0][0] = c + t.x * a.x; r[0][1] = t.x * a.y + s * a.z; r[0][2] = t.x * a.z - s * a.y;
  r[1][0] = t.y * a.x - s * a.z; r[1][1] = c + t.y * a.y; r[1][2] = t.y * a.z + s * a.x;
  r[2][0] = t.z * a.x + s * a.y; r[2][1] = t.z * a.y - s * a.x; r[2][2] = c + t.z * a.z;
  Mat4 out{};
  for (int k = 0; k < 3; ++k)
    for (int i = 0; i < 4; ++i) out.m[k][i] = m.m[0][i] * r[k][0] + m.m[1][i] * r[k][1] + m.m[2][i] * r[k][2];
  for (int i = 0; i < 4; ++i) out.m[3][i] = m.m[3][i];
  return out;
}

// glm a * b (type_mat4x4.inl:630-648), left-to-right sums
[[maybe_unused]] Mat4 mat_mul(const Mat4 &a, const Mat4 &b) {
  Mat4 r{};
  for (int c = 0; c < 4; ++c)
    for (int k = 0; k < 4; ++k)
      r.m[c][k] = ((a.m[0][k] * b.m[c][0] + a.m[1][k] * b.m[c][1]) + a.m[2][k] * b.m[c][2]) + a.m[3][k] * b.m[c][3];
  return r;
}

// glm::translate (ext/matrix_transform.inl:10-15)
[[maybe_unused]] Mat4 translate(const Mat4 &m, Vec3 v) {
  Mat4 r = m;
  for (int k = 0; k < 4; ++k) r.m[3][k] = m.m[0][k] * v.x + m.m[1][k] * v.y + m.m[2][k] * v.z + m.m[3][k];
  return r;
}

Vec3 transform_point(const Mat4 &m, Vec3 p) {   // glm mat4 * vec4(p, 1), xyz
  float o[3];
  for (int i = 0; i < 3; ++i) o[i] = (m.m[0][i] * p.x + m.m[1][i] * p.y) + (m.m[2][i] * p.z + m.m[3][i] * 1.f);
  return {o[0], o[1], o[2]};
}

// src/main/splat.cpp:235-241
Mat4 mvp_start() {
  Mat4 m = identity();
  m.m[0][0] = -1.f;
  m.m[1][1] = -0.09709989f; m.m[1][2] = -0.99527466f;
  m.m[2][1] = -0.99527466f; m.m[2][2] = 0.09709989f;
  m.m[3][2] = -5.1539507f;
  return m;
}

// src/main/splat.cpp:186-195 with camera.cpp:10-15 and geometry.cpp:9-24
Mat4 app_projection(Vec3 bmin, Vec3 bmax, float fov, float aspect) {
  const Vec3 centre = {(bmax.x + bmin.x) * .5f, (bmax.y + bmin.y) * .5f, (bmax.z + bmin.z) * .5f};
  const Vec3 d = sub(bmax, bmin);
  const float radius = std::sqrt(dot(d, d)) * .5f;
  const Mat4 mv = look_at({centre.x, centre.y, centre.z - 1.f * radius}, centre, {0.f, 1.f, 1.f});
  const Vec3 cmin = transform_point(mv, bmin), cmax = transform_point(mv, bmax), cd = sub(cmax, cmin);
  const float r = std::sqrt(dot(cd, cd)) * .5f;
  const float nr = r / std::tan(fov), fr = nr + 20.f * r;
  return frustum(-r * aspect, r * aspect, -r, r, nr, fr);
}

// Binary-little-endian PLY with float vertex properties (what happly reads for the reference,
// src/splat/file_io.cpp:30-42,62-77).  Returns one column per requested property.
[[maybe_unused]] bool read_ply(const std::string &path, const std::vector<std::string> &want,
                               std::vector<std::vector<float>> &cols, std::size_t &n, std::string &err) {
  std::ifstream fh(path, std::ios::binary);
  if (!fh) { err = "cannot open " + path; return false; }
  std::vector<std::string> names;
  std::string line, fmt;
  n = 0;
  bool in_vertex = false;
  while (std::getline(fh, line)) {
    std::istringstream ss(line);
    std::string tok;
    ss >> tok;
    if (tok == "format") ss >> fmt;
    else if (tok == "element") {
      std::string what;
      ss >> what;
      in_vertex = what == "vertex";
      if (in_vertex) ss >> n;
    } else if (tok == "property" && in_vertex) {
      std::string ty, nm;
      ss >> ty >> nm;
      if (ty != "float") { err = "non-float vertex property " + nm; return false; }
      names.push_back(nm);
    } else if (tok == "end_header") break;
  }
  if (fmt != "binary_little_endian") { err = "unsupported PLY format '" + fmt + "'"; return false; }
  std::vector<float> raw(n * names.size());
  fh.read(reinterpret_cast<char *>(raw.data()), (std::streamsize)(raw.size() * sizeof(float)));
  if ((std::size_t)fh.gcount() != raw.size() * sizeof(float)) { err = "truncated vertex data"; return false; }
  cols.assign(want.size(), std::vector<float>(n));
  for (std::size_t w = 0; w < want.size(); ++w) {
    const auto it = std::find(names.begin(), names.end(), want[w]);
    if (it == names.end()) { err = "missing property " + want[w]; return false; }
    const std::size_t c = (std::size_t)(it - names.begin());
    for (std::size_t i = 0; i < n; ++i) cols[w][i] = raw[i * names.size() + c];
  }
  return true;
}

#ifdef GSPLAT_WITH_HAPPLY
// The reference's loader: happly::PLYData + getElement("vertex").getProperty<float>(name) for the 14
// properties, exactly the calls of fillProperty / fillPlyProperties (src/splat/file_io.cpp:57-77).
bool read_ply_happly(const std::string &path, const std::vector<std::string> &want, std::vector<std::vector<float>> &cols,
                     std::size_t &n, std::string &err) {
  try {
    happly::PLYData plyData(path);
    cols.clear();
    for (const auto &name : want) cols.push_back(plyData.getElement("vertex").getProperty<float>(name));
    n = cols.empty() ? 0 : cols[0].size();
  } catch (const std::exception &e) {
    err = e.what();
    return false;
  }
  return true;
}
#endif

// src/main/splat.cpp:136-163 + :93-100
bool gaussians_from_ply(const std::string &path, std::vector<splat::Gaussian3DRecord> &g, std::string &err) {
  const std::vector<std::string> want = {"x", "y", "z", "f_dc_0", "f_dc_1", "f_dc_2", "opacity", "scale_0", "scale_1",
                                         "scale_2", "rot_0", "rot_1", "rot_2", "rot_3"};
  std::vector<std::vector<float>> c;
  std::size_t n = 0;
#ifdef GSPLAT_WITH_HAPPLY
  if (!read_ply_happly(path, want, c, n, err)) return false;
#else
  if (!read_ply(path, want, c, n, err)) return false;
#endif
  g.resize(n);
  const float SH_C0 = 0.28209479177387814f;
  for (std::size_t i = 0; i < n; ++i) {
    splat::Gaussian3DRecord &r = g[i];
    r.mean[0] = c[0][i]; r.mean[1] = c[1][i]; r.mean[2] = c[2][i]; r.mean[3] = 1.f;
    for (int k = 0; k < 3; ++k) r.colour[k] = std::max(SH_C0 * c[3 + k][i] + 0.5f, 0.f);   // no upper clamp
    r.colour[3] = c[6][i];                                                                // raw (logit) opacity
    for (int k = 0; k < 3; ++k) r.scale[k] = c[7 + k][i];                                  // raw log-scales
    for (int k = 0; k < 4; ++k) r.rot[k] = c[10 + k][i];                                   // rot_0 = quaternion w
    r.gid = (float)(i + 1);
  }
  return true;
}

void recentre_negate_z(std::vector<splat::Gaussian3DRecord> &g, Vec3 &bmin, Vec3 &bmax) {
  bmin = {1e30f, 1e30f, 1e30f};
  bmax = {-1e30f, -1e30f, -1e30f};
  for (const auto &r : g) {
    bmin = {std::min(bmin.x, r.mean[0]), std::min(bmin.y, r.mean[1]), std::min(bmin.z, r.mean[2])};
    bmax = {std::max(bmax.x, r.mean[0]), std::max(bmax.y, r.mean[1]), std::max(bmax.z, r.mean[2])};
  }
  const Vec3 centre = {(bmax.x + bmin.x) * .5f, (bmax.y + bmin.y) * .5f, (bmax.z + bmin.z) * .5f};
  bmin = {1e30f, 1e30f, 1e30f};
  bmax = {-1e30f, -1e30f, -1e30f};
  for (auto &r : g) {
    r.mean[0] -= centre.x; r.mean[1] -= centre.y; r.mean[2] = -(r.mean[2] - centre.z);
    bmin = {std::min(bmin.x, r.mean[0]), std::min(bmin.y, r.mean[1]), std::min(bmin.z, r.mean[2])};
    bmax = {std::max(bmax.x, r.mean[0]), std::max(bmax.y, r.mean[1]), std::max(bmax.z, r.mean[2])};
  }
}

const char *arg(int argc, char **argv, const char *name, const char *dflt) {
  for (int i = 1; i + 1 < argc; ++i)
    if (!std::strcmp(argv[i], name)) return argv[i + 1];
  return dflt;
}

}  // namespace

// ---- scene cache ---------------------------------------------------------------
// The converted scene — Gaussian3D records after the PLY -> record conversion and the app's
// recentre + z-negate — as one flat file, so that a restart skips PLY parsing and conversion:
//   "GSPLB200" | u32 version (1) | u32 0 | u64 n | f32 bbox min[3], max[3] | n x 16 f32 records
// (little-endian; gaussian_splat_ipu_b200/scene.py reads and writes the same file).
static bool cache_write(const std::string &path, const std::vector<splat::Gaussian3DRecord> &g, Vec3 bmin, Vec3 bmax) {
  std::FILE *f = std::fopen(path.c_str(), "wb");
  if (!f) return false;
  const std::uint32_t ver[2] = {1u, 0u};
  const std::uint64_t n = g.size();
  const float bb[6] = {bmin.x, bmin.y, bmin.z, bmax.x, bmax.y, bmax.z};
  bool ok = std::fwrite("GSPLB200", 1, 8, f) == 8 && std::fwrite(ver, 4, 2, f) == 2 && std::fwrite(&n, 8, 1, f) == 1 &&
            std::fwrite(bb, 4, 6, f) == 6 && std::fwrite(g.data(), sizeof(splat::Gaussian3DRecord), g.size(), f) == g.size();
  return std::fclose(f) == 0 && ok;
}
static bool cache_read(const std::string &path, std::vector<splat::Gaussian3DRecord> &g, Vec3 &bmin, Vec3 &bmax) {
  std::FILE *f = std::fopen(path.c_str(), "rb");
  if (!f) return false;
  char magic[8];
  std::uint32_t ver[2];
  std::uint64_t n = 0;
  float bb[6];
  bool ok = std::fread(magic, 1, 8, f) == 8 && !std::memcmp(magic, "GSPLB200", 8) && std::fread(ver, 4, 2, f) == 2 &&
            ver[0] == 1u && std::fread(&n, 8, 1, f) == 1 && std::fread(bb, 4, 6, f) == 6;
  if (ok) {
    g.resize((std::size_t)n);
    ok = std::fread(g.data(), sizeof(splat::Gaussian3DRecord), g.size(), f) == g.size();
    bmin = {bb[0], bb[1], bb[2]}; bmax = {bb[3], bb[4], bb[5]};
  }
  std::fclose(f);
  return ok;
}

int main(int argc, char **argv) {
  const std::string input = arg(argc, argv, "--input", "");
  const std::size_t synthetic = std::strtoull(arg(argc, argv, "--synthetic", "0"), nullptr, 10);
  const int W = std::atoi(arg(argc, argv, "--width", "1280")), H = std::atoi(arg(argc, argv, "--height", "720"));
  const int frames = std::max(1, std::atoi(arg(argc, argv, "--frames", "32")));
  const float scale_div = (float)std::atof(arg(argc, argv, "--scale-div", synthetic ? "1.0" : "0.1"));   // lambda1/10, splat.cpp:262
  const std::string out = arg(argc, argv, "--out", "");
  const int device = std::atoi(arg(argc, argv, "--device", "0"));
  const std::string cache = arg(argc, argv, "--cache", "");   // scene cache: read if present, else written after the PLY conversion
  bool cache_only = false;                                     // stop after the scene is built / cached (no device needed)
  for (int i = 1; i < argc; ++i) cache_only = cache_only || !std::strcmp(argv[i], "--cache-only");
  if (input.empty() && !synthetic && cache.empty()) {
    std::fprintf(stderr, "usage: %s --input scene.ply | --synthetic N  [--width W --height H --frames K --scale-div D "
                         "--out frame.ppm --device I --cache scene.gsc]\n", argv[0]);
    return 2;
  }
  std::vector<splat::Gaussian3DRecord> g;
  Vec3 bmin{}, bmax{};
  if (synthetic) {
    g.resize(synthetic);
    if (gsplat_synth_scene(0xB2000000ull, synthetic, reinterpret_cast<float *>(g.data())) != GSPLAT_OK) return 2;
    Vec3 lo = {1e30f, 1e30f, 1e30f}, hi = {-1e30f, -1e30f, -1e30f};
    for (const auto &r : g) {   // the generator has already recentred and negated z
      lo = {std::min(lo.x, r.mean[0]), std::min(lo.y, r.mean[1]), std::min(lo.z, r.mean[2])};
      hi = {std::max(hi.x, r.mean[0]), std::max(hi.y, r.mean[1]), std::max(hi.z, r.mean[2])};
    }
    bmin = lo; bmax = hi;
  } else if (!cache.empty() && cache_read(cache, g, bmin, bmax)) {
    std::fprintf(stderr, "scene cache %s: %zu Gaussians\n", cache.c_str(), g.size());
  } else {
    std::string err;
    if (!gaussians_from_ply(input, g, err)) { std::fprintf(stderr, "%s: %s\n", input.c_str(), err.c_str()); return 2; }
    recentre_negate_z(g, bmin, bmax);
    if (!cache.empty() && !cache_write(cache, g, bmin, bmax)) std::fprintf(stderr, "cannot write scene cache %s\n", cache.c_str());
  }
  if (cache_only) {
    std::printf("ok gaussians %zu cache %s\n", g.size(), cache.c_str());
    return 0;
  }
  const float fov = 40.0f * 0.017453292519943295f;   // glm::radians(40.f), splat.cpp:170
  const Mat4 proj = app_projection(bmin, bmax, fov, (float)W / (float)H);
  const Mat4 base = mvp_start();
  const int ui_port = std::atoi(arg(argc, argv, "--ui-port", "0"));
#ifdef GSPLAT_WITH_UI
  if (ui_port) {
    // The interactive loop of src/main/splat.cpp:245-322.  uiServer->start() blocks until the viewer has
    // connected (InterfaceServer.hpp:272-278); per frame: the UI state gives the view
    // (modelView * Rx(env) * Ry(env2) * T(X/50, Y/50, -Z/20 + 20), :312-314), the projection
    // (fitFrustumToBoundingBox(bbInCamera, state.fov, aspect), :287) and fxy = (fov, lambda1 / 10) (:262);
    // the finished frame and the tile histogram go to the viewer while the next frame renders (the
    // AsyncTask + imagePtr / imagePtrBuffered swap, :208-225,281-283 — here the frame stream's slots).
    try {
      splat::B200Splatter s(g.data(), g.size(), W, H, device);
      b200ui::B200StreamServer ui(ui_port);
      ui.start();
      if (!ui.connected()) { std::printf("ui: no client\n"); return 4; }
      ui.initialiseVideoStream(W, H);
      b200ui::B200StreamServer::State state;
      state.fov = fov;                               // splat.cpp:175
      ui.updateFov(state.fov);
      const Vec3 centre = {(bmax.x + bmin.x) * .5f, (bmax.y + bmin.y) * .5f, (bmax.z + bmin.z) * .5f};
      const Vec3 d = sub(bmax, bmin);
      const float radius = std::sqrt(dot(d, d)) * .5f;
      const Mat4 modelView = look_at({centre.x, centre.y, centre.z - 1.f * radius}, centre, {0.f, 1.f, 1.f});   // :186
      Mat4 dynamicView = base, projection = proj;
      std::vector<std::uint64_t> inflight;
      const std::size_t depth = (std::size_t)s.concurrentFrames();
      std::vector<std::uint32_t> hist((std::size_t)s.numTiles());
      unsigned long long frames_sent = 0;
      auto deliver = [&](std::uint64_t ticket) {
        TraceScope trace("ui_update");
        const std::uint8_t *bgr = nullptr;
        const std::uint32_t *counts = nullptr;
        s.acquireFrame(ticket, &bgr, &counts);
        hist.assign(counts, counts + s.numTiles());
        ui.sendHistogram(hist);
        ui.sendPreviewImage(bgr, W, H, (std::size_t)W * 3);
        ++frames_sent;
      };
      do {
        {
          TraceScope trace("mvp_transform_ipu");
          s.updateModelView(dynamicView);
          s.updateProjection(projection);
          s.updateFocalLengths(state.fov, state.lambda1 / 10.f);
          inflight.push_back(s.executeStreamed());
        }
        if (inflight.size() > depth) { deliver(inflight.front()); inflight.erase(inflight.begin()); }
        state = ui.consumeState();
        projection = app_projection(bmin, bmax, state.fov, (float)W / (float)H);
        const float rad = 0.017453292519943295f;
        dynamicView = mat_mul(modelView, rotate(identity(), state.envRotationDegrees * rad, {1.f, 0.f, 0.f}));
        dynamicView = rotate(dynamicView, state.envRotationDegrees2 * rad, {0.f, 1.f, 0.f});
        dynamicView = translate(dynamicView, {state.X / 50.f, state.Y / 50.f, -state.Z / 20.f + 20.f});
        // a viewer that reads slower than the GPU renders must not grow the muxer's queue without bound
        std::this_thread::sleep_for(std::chrono::microseconds(500));
      } while (!state.stop && ui.connected());
      while (!inflight.empty() && ui.connected()) { deliver(inflight.front()); inflight.erase(inflight.begin()); }
      std::printf("ok ui frames %llu\n", frames_sent);
    } catch (const std::runtime_error &e) {
      std::printf("no-device: %s\n", e.what());
      return 3;
    }
    return 0;
  }
#else
  if (ui_port) { std::fprintf(stderr, "--ui-port: built without packetcomms (reference tree not reachable at build time)\n"); return 2; }
#endif
  try {
    splat::B200Splatter s(g.data(), g.size(), W, H, device);
    s.updateProjection(proj);
    s.updateFocalLengths(fov, scale_div);
    const std::uint8_t *bgr = nullptr;
    const std::uint32_t *counts = nullptr;
    std::vector<std::uint64_t> inflight;
    const std::size_t depth = 2 * (std::size_t)s.concurrentFrames();   // tickets outstanding: two per frame in flight
    unsigned long long pixel_sum = 0, count_sum = 0;
    auto deliver = [&](std::uint64_t ticket) {   // what the AsyncTask thread does with imagePtrBuffered
      TraceScope trace("ui_update");
      s.acquireFrame(ticket, &bgr, &counts);
      pixel_sum = 0; count_sum = 0;
      for (std::size_t i = 0; i < (std::size_t)W * H * 3; i += 97) pixel_sum += bgr[i];
      for (int t = 0; t < s.numTiles(); ++t) count_sum += counts[t];
    };
    const auto t0 = std::chrono::steady_clock::now();
    for (int k = 0; k < frames; ++k) {
      // config 2 orbit: glm::rotate(mvpStart, radians(360 / frames * k), (0,1,0))
      TraceScope trace("mvp_transform_ipu");
      s.updateModelView(rotate(base, (360.0f / (float)frames * (float)k) * 0.017453292519943295f, {0.f, 1.f, 0.f}));
      inflight.push_back(s.executeStreamed());
      if (inflight.size() > depth) { deliver(inflight.front()); inflight.erase(inflight.begin()); }
    }
    while (!inflight.empty()) { deliver(inflight.front()); inflight.erase(inflight.begin()); }
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    const gsplat_frame_stats st = s.stats();
    if (!out.empty()) {   // last frame, BGR -> RGB
      std::ofstream fh(out, std::ios::binary);
      fh << "P6\n" << W << " " << H << "\n255\n";
      std::vector<std::uint8_t> row((std::size_t)W * 3);
      for (int y = 0; y < H; ++y) {
        for (int x = 0; x < W; ++x) {
          const std::uint8_t *p = bgr + ((std::size_t)y * W + x) * 3;
          row[x * 3 + 0] = p[2]; row[x * 3 + 1] = p[1]; row[x * 3 + 2] = p[0];
        }
        fh.write(reinterpret_cast<const char *>(row.data()), (std::streamsize)row.size());
      }
    }
    std::printf("ok gaussians %zu frames %d fps %.1f drawn %llu pairs %llu count_sum %llu pixel_sum %llu\n", g.size(), frames,
                frames / sec, (unsigned long long)st.num_drawn, (unsigned long long)st.num_reference_duplicates, count_sum,
                pixel_sum);
  } catch (const std::runtime_error &e) {
    std::printf("no-device: %s\n", e.what());
    return 3;
  }
  return 0;
}
