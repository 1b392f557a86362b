// oracle/ref_driver.cpp — TEST INFRASTRUCTURE, not product code.
//
// Thin C-ABI driver around the UNMODIFIED reference splat code.  build_ref.sh pipes this file
// to g++ *after* the reference's own sources (read where they lie under /root/reference):
//   include/math/sincos.hpp, include/splat/ipu_geometry.hpp, include/tileMapping/tile_config.hpp,
//   include/splat/viewport.hpp, codelets/splat/codelets.cpp:1-642
// so every arithmetic step below is executed by the reference's own functions:
//   Viewport::clipSpaceToViewport   include/splat/viewport.hpp:21-35
//   Gaussian3D::ComputeCov2D        include/splat/ipu_geometry.hpp:333-383
//   Gaussian2D::GetBoundingBox      include/splat/ipu_geometry.hpp:263-276
//   Bounds2f::contains / diagonal   include/splat/ipu_geometry.hpp:113-165
//   TiledFramebuffer::getTileBounds include/tileMapping/tile_config.hpp:57-71
//   GSplat::renderTile (quicksort + ComputeConicOpacity + blend)  codelets/splat/codelets.cpp:303-421
// The only logic restated here is the converged-lattice tile rectangle and the
// anchoring predicate (a6) — i.e. WHICH tiles hold a Gaussian once the 4-neighbour lattice of
// src/splat/edge_builder.cpp:35-84 has converged with capacities lifted.  lattice (O1) mode runs
// GSplat::compute itself over the lattice to prove that restatement (see ref_lattice_*).
//
// The image geometry macros of tile_config.hpp:5-8 are rewritten by build_ref.sh to the
// runtime variables below so one binary serves 1280x720, 1920x1080 and 3840x2160.

#include <algorithm>
#include <vector>
#include <chrono>
#include <omp.h>

// NOTE: g_ref_imwidth etc. are declared in the prologue build_ref.sh emits before the reference
// headers (they must be visible where the macros are used).

namespace {

struct Frame {
  int W = 0, H = 0, TA = 0, TD = 0;
  std::size_t n = 0;
  std::vector<float> rec;        // n x 8: px, py, clip_z, cov_a, cov_b, cov_c, radius, drawn(0/1)
  std::vector<int32_t> rect;     // n x 4: c0, c1, r0, r1 (tile units, clamped), valid iff drawn
  std::vector<uint32_t> counts;  // numTiles
  std::vector<int64_t> offsets;  // numTiles + 1
  std::vector<int32_t> lists;    // D entries, gaussian index (gid-1), canonical (z asc, gid asc)
  std::vector<float> fb;         // W*H*4, tile-major chunks of tw*th*4 (ipu_rasteriser.cpp:280-285)
  double project_ms = 0, bin_ms = 0, blend_ms = 0;
  uint64_t pairs = 0;
};

double now_ms() {
  using namespace std::chrono;
  return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

void set_geometry(int W, int H, int TA, int TD) {
  g_ref_imwidth = float(W);
  g_ref_imheight = float(H);
  g_ref_tiles_across = float(TA);
  g_ref_tiles_down = float(TD);
}

} // namespace

extern "C" {

// Projects every Gaussian exactly as GSplat::renderInternal does (codelets.cpp:437-470) and
// derives the render predicate + converged tile rectangle.
// g16: n records of 16 floats (the reference's Gaussian3D: mean4 colour4 rot4 scale3 gid).
// mv/proj: glm column-major 4x4 (what IpuSplatter::updateModelView receives).
// 0: ties in clip z blend in (z, gid) order (the parity contract); 1: the reference's quicksort decides
static int g_tie_order = 0;
void ref_set_tie_order(int quicksort_decides) { g_tie_order = quicksort_decides ? 1 : 0; }

void *ref_frame_create(const float *g16, std::size_t n, const float *mv, const float *proj,
                       float fov, float scale_div, int W, int H, int TA, int TD, int do_render,
                       int nthreads) {
  set_geometry(W, H, TA, TD);
  auto *f = new Frame;
  f->W = W; f->H = H; f->TA = TA; f->TD = TD; f->n = n;
  f->rec.assign(n * 8, 0.f);
  f->rect.assign(n * 4, 0);
  const int numTiles = TA * TD;
  if (nthreads <= 0) nthreads = omp_get_max_threads();

  // Matrix path of the codelet: host uploads transpose (ipu_rasteriser.cpp:86-102), vertex
  // transposes back (codelets.cpp:625-628) => the original glm matrices.
  const glm::mat4 viewT = glm::transpose(glm::make_mat4(mv));
  const glm::mat4 projT = glm::transpose(glm::make_mat4(proj));
  const glm::mat4 viewmatrix = glm::transpose(glm::make_mat4(glm::value_ptr(viewT)));
  const glm::mat4 projmatrix = glm::transpose(glm::make_mat4(glm::value_ptr(projT)));

  const TiledFramebuffer tfb(IPU_TILEWIDTH, IPU_TILEHEIGHT);
  const splat::Viewport vp(0.0f, 0.0f, IMWIDTH, IMHEIGHT);
  const float clipSize = 15.0f;                       // codelets.cpp:622
  const float fxy[2] = {fov, scale_div};
  const auto tb0 = tfb.getTileBounds(0);
  const Bounds2f image({0.f, 0.f}, {float(W), float(H)});
  const float tw = tb0.max.x - tb0.min.x, th = tb0.max.y - tb0.min.y;

  double t0 = now_ms();
  // codelets.cpp:443-448
  const auto mvp = projmatrix * viewmatrix;
  glm::vec2 tanfov(tan(0.5 * fxy[0]), tan(0.5 * fxy[0]));
  glm::vec2 focal(tfb.width / (2.f * glm::tan(fxy[0] / 2.f)),
                  tfb.height / (2.f * glm::tan(fxy[0] / 2.f)));

#pragma omp parallel for num_threads(nthreads) schedule(static)
  for (std::size_t i = 0; i < n; ++i) {
    Gaussian3D g;
    std::memcpy(&g, g16 + i * 16, sizeof(g));
    float *r = &f->rec[i * 8];
    int32_t *rc = &f->rect[i * 4];
    rc[0] = 0; rc[1] = -1; rc[2] = 0; rc[3] = -1;
    if (g.gid <= 0) continue;                         // codelets.cpp:456
    auto clipSpace = mvp * glm::vec4(g.mean.x, g.mean.y, g.mean.z, g.mean.w);   // :460
    auto projMean = vp.clipSpaceToViewport(clipSpace);                           // :461
    g.scale = g.scale / fxy[1];                                                  // :463
    ivec3 cov2D = g.ComputeCov2D(projmatrix, viewmatrix, tanfov.x, tanfov.y, focal.x, focal.y);
    Gaussian2D g2D({projMean.x, projMean.y}, g.colour, cov2D, clipSpace.z);      // :466
    auto bb = g2D.GetBoundingBox();                                              // :467
    bool withinGuardBand = bb.diagonal().length() < tb0.diagonal().length() * clipSize;  // :470
    r[0] = projMean.x; r[1] = projMean.y; r[2] = clipSpace.z;
    r[3] = cov2D.x; r[4] = cov2D.y; r[5] = cov2D.z;
    r[6] = (bb.max.x - bb.min.x) * 0.5f;   // informational only
    // a6: drawn iff guard band, z<0 (codelets.cpp:493) and the mean is anchored by some tile
    // (Bounds2f::contains over the union of tiles == the image rectangle).
    bool drawn = withinGuardBand && (g2D.z < 0.0f) && image.contains(g2D.mean);
    r[7] = drawn ? 1.f : 0.f;
    if (!drawn) continue;
    // a7: converged lattice membership == rectangle of tiles reached by Bounds2f::clip's
    // propagation rule (left while floor(min.x) < tile.min.x; right while ceil(max.x) >= tile.max.x).
    int c0 = int(std::floor(floor(bb.min.x) / tw));
    int c1 = int(std::floor(ceil(bb.max.x) / tw));
    int r0 = int(std::floor(floor(bb.min.y) / th));
    int r1 = int(std::floor(ceil(bb.max.y) / th));
    rc[0] = std::max(c0, 0); rc[1] = std::min(c1, TA - 1);
    rc[2] = std::max(r0, 0); rc[3] = std::min(r1, TD - 1);
  }
  f->project_ms = now_ms() - t0;

  // Binning (counting sort by tile, gid order inside a tile).
  t0 = now_ms();
  f->counts.assign(numTiles, 0);
  for (std::size_t i = 0; i < n; ++i) {
    if (f->rec[i * 8 + 7] == 0.f) continue;
    const int32_t *rc = &f->rect[i * 4];
    for (int ty = rc[2]; ty <= rc[3]; ++ty)
      for (int tx = rc[0]; tx <= rc[1]; ++tx) f->counts[ty * TA + tx]++;
  }
  f->offsets.assign(numTiles + 1, 0);
  for (int t = 0; t < numTiles; ++t) f->offsets[t + 1] = f->offsets[t] + f->counts[t];
  f->lists.assign(f->offsets[numTiles], 0);
  {
    std::vector<int64_t> cur(f->offsets.begin(), f->offsets.end() - 1);
    for (std::size_t i = 0; i < n; ++i) {
      if (f->rec[i * 8 + 7] == 0.f) continue;
      const int32_t *rc = &f->rect[i * 4];
      for (int ty = rc[2]; ty <= rc[3]; ++ty)
        for (int tx = rc[0]; tx <= rc[1]; ++tx) f->lists[cur[ty * TA + tx]++] = int32_t(i);
    }
  }
  f->bin_ms = now_ms() - t0;

  if (!do_render) {
    // canonical order for list comparison even without rendering
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 8)
    for (int t = 0; t < numTiles; ++t) {
      auto b = f->lists.begin() + f->offsets[t], e = f->lists.begin() + f->offsets[t + 1];
      std::stable_sort(b, e, [&](int32_t a, int32_t c) { return f->rec[a * 8 + 2] < f->rec[c * 8 + 2]; });
    }
    return f;
  }

  // Blend: per tile, fill a GSplat vertex's gaus2D and call the reference's own renderTile
  // (ComputeConicOpacity and the blend loop run unmodified).
  //
  // Tie order.  Different Gaussians do share a clip-space z bit for bit (about 150 pairs per frame on
  // the shipped PLYs), the reference's quicksort is unstable (codelets.cpp:303-344), and on the IPU
  // the buffer it sorts is in lattice ARRIVAL order, which changes from run to run: the order of tied
  // entries is unspecified in the reference.  It is not harmless either: swapping two tied entries
  // moves T by an ulp, and pixels that sit on the T * (1 - alpha) < 1e-4 stop (stacks of opaque
  // splats: 1 -> 0.01 -> 1e-4) then flip by up to colour * 0.99 * 0.01.  The parity contract
  //  fixes the order of ties as (z ascending, gid ascending):
  //   g_tie_order == 0 (default): the list is put in that order BEFORE it is handed to the vertex,
  //     and `indices` is bound with a single element, for which the reference's own sortBuffer
  //     returns without sorting (codelets.cpp:348-350: end >= indices.size() - 1) — an ordered list
  //     is also the O(n^2) case of its last-element-pivot quicksort (minutes on a 363k-entry tile);
  //   g_tie_order == 1: gid-order fill, the reference's quicksort runs and decides the ties (what
  //     round 1 did; kept to measure how far the two orders are apart).
  t0 = now_ms();
  f->fb.assign(std::size_t(W) * H * 4, 0.f);
  const std::size_t tilePix = std::size_t(tw) * std::size_t(th);
  const std::size_t g2stride = sizeof(Gaussian2D);   // the codelet strides by sizeof in FLOATS (codelets.cpp:370,494)
  const bool canonical = g_tie_order == 0;
#pragma omp parallel num_threads(nthreads)
  {
    std::vector<float> gaus2D;
    std::vector<int> indices;
    int tileId[1];
    GSplat v;
#pragma omp for schedule(dynamic, 1)
    for (int t = 0; t < numTiles; ++t) {
      const std::size_t cnt = f->counts[t];
      float *fbTile = &f->fb[std::size_t(t) * tilePix * 4];
      if (cnt == 0) continue;                          // codelets.cpp:501 (fb stays zero: colourFb)
      auto b = f->lists.begin() + f->offsets[t], e = f->lists.begin() + f->offsets[t + 1];
      auto by_z = [&](int32_t a, int32_t c) { return f->rec[a * 8 + 2] < f->rec[c * 8 + 2]; };
      if (canonical) std::stable_sort(b, e, by_z);     // (z asc, gid asc): the list was in gid order
      gaus2D.assign((cnt + 2) * g2stride, 0.f);        // slot n is the all-zero sentinel (a8)
      indices.assign(canonical ? 1 : 2 * cnt + 64, 0);
      for (std::size_t k = 0; k < cnt; ++k) {
        const std::size_t i = f->lists[f->offsets[t] + k];
        Gaussian3D g;
        std::memcpy(&g, g16 + i * 16, sizeof(g));
        const float *r = &f->rec[i * 8];
        Gaussian2D g2D({r[0], r[1]}, g.colour, {r[3], r[4], r[5]}, r[2]);
        std::memcpy(&gaus2D[k * g2stride], &g2D, sizeof(g2D));   // insertAt, codelets.cpp:494-495
      }
      tileId[0] = t;
      v.gaus2D.bind(gaus2D.data(), gaus2D.size());
      v.indices.bind(indices.data(), indices.size());
      v.localFb.bind(fbTile, tilePix * 4);
      v.tile_id.bind(tileId, 1);
      v.renderTile(cnt, tfb.getTileBounds(t));
      // the lists are reported in the canonical order in both modes
      if (!canonical) std::stable_sort(b, e, by_z);
    }
  }
  f->blend_ms = now_ms() - t0;
  return f;
}

void ref_frame_destroy(void *h) { delete static_cast<Frame *>(h); }
std::size_t ref_frame_num_duplicates(void *h) { return static_cast<Frame *>(h)->lists.size(); }
void ref_frame_get_records(void *h, float *rec8, int32_t *rect4) {
  auto *f = static_cast<Frame *>(h);
  std::memcpy(rec8, f->rec.data(), f->rec.size() * sizeof(float));
  std::memcpy(rect4, f->rect.data(), f->rect.size() * sizeof(int32_t));
}
void ref_frame_get_lists(void *h, uint32_t *counts, int64_t *offsets, int32_t *lists) {
  auto *f = static_cast<Frame *>(h);
  std::memcpy(counts, f->counts.data(), f->counts.size() * sizeof(uint32_t));
  std::memcpy(offsets, f->offsets.data(), f->offsets.size() * sizeof(int64_t));
  if (!f->lists.empty()) std::memcpy(lists, f->lists.data(), f->lists.size() * sizeof(int32_t));
}
void ref_frame_get_framebuffer(void *h, float *fb) {
  auto *f = static_cast<Frame *>(h);
  if (!f->fb.empty()) std::memcpy(fb, f->fb.data(), f->fb.size() * sizeof(float));
}
void ref_frame_get_timings(void *h, double *ms3) {
  auto *f = static_cast<Frame *>(h);
  ms3[0] = f->project_ms; ms3[1] = f->bin_ms; ms3[2] = f->blend_ms;
}
int ref_num_threads(void) { return omp_get_max_threads(); }

// Known-answer helpers: single reference functions, for per-function KATs.
void ref_conic_opacity(const float *cov3, float opacity, float *out4) {
  Gaussian2D g({0.f, 0.f}, {0.f, 0.f, 0.f, opacity}, {cov3[0], cov3[1], cov3[2]}, -1.f);
  auto c = g.ComputeConicOpacity();                     // ipu_geometry.hpp:278-286
  out4[0] = c.x; out4[1] = c.y; out4[2] = c.z; out4[3] = c.w;
}
void ref_tile_bounds(int W, int H, int TA, int TD, unsigned tid, float *out4) {
  set_geometry(W, H, TA, TD);
  const TiledFramebuffer tfb(IPU_TILEWIDTH, IPU_TILEHEIGHT);
  auto b = tfb.getTileBounds(tid);                      // tile_config.hpp:57-71
  out4[0] = b.min.x; out4[1] = b.min.y; out4[2] = b.max.x; out4[3] = b.max.y;
}

// ---------------------------------------------------------------------------------------------
// O1: the reference lattice itself.  numTiles GSplat vertices + the 4-neighbour channels of
// src/splat/edge_builder.cpp:35-84 and the buffer sizing of src/splat/ipu_rasteriser.cpp:291-366,
// with the two capacities (spare slots per tile, Gaussians per channel) as parameters so the
// as-shipped (600 / 75... see ipu_rasteriser.cpp:307-309) and the lifted configurations both run.
// ---------------------------------------------------------------------------------------------
struct Lattice {
  int W, H, TA, TD, numTiles;
  std::size_t slotsPerTile, chanSlots;
  std::vector<GSplat> verts;
  std::vector<std::vector<float>> vertsIn, gaus2D, fb;
  std::vector<std::vector<int>> indices;
  std::vector<std::vector<float>> out[4], in[4];   // order: right, left, up, down
  std::vector<std::array<int, 1>> tid;
  std::vector<std::array<unsigned, 1>> splatted;
  std::vector<float> mvT, projT, fxy;
  struct Edge { int src, srcDir, dst, dstDir; };
  std::vector<Edge> edges;
  double compute_ms = 0, exchange_ms = 0;
};

void *ref_lattice_create(const float *g16, std::size_t n, int W, int H, int TA, int TD,
                         std::size_t spareSlots, std::size_t chanSlots) {
  set_geometry(W, H, TA, TD);
  auto *L = new Lattice;
  L->W = W; L->H = H; L->TA = TA; L->TD = TD; L->numTiles = TA * TD;
  const int T = L->numTiles;
  const std::size_t S = sizeof(Gaussian3D);          // stride in FLOATS (codelets.cpp:452)
  // even split in file order (calculateMapping/applyTileMapping, ipu_rasteriser.cpp:164-214)
  const std::size_t perTile = (n + T - 1) / T;
  L->slotsPerTile = perTile + spareSlots;
  L->chanSlots = chanSlots;
  L->verts.resize(T);
  L->vertsIn.assign(T, std::vector<float>(L->slotsPerTile * S, 0.f));
  L->gaus2D.assign(T, std::vector<float>(L->slotsPerTile * sizeof(Gaussian2D), 0.f));
  L->indices.assign(T, std::vector<int>(L->slotsPerTile, 0));
  L->fb.assign(T, std::vector<float>(std::size_t(W / TA) * (H / TD) * 4, 0.f));
  for (int d = 0; d < 4; ++d) {
    L->out[d].assign(T, std::vector<float>(chanSlots * S, 0.f));
    L->in[d].assign(T, std::vector<float>(chanSlots * S, 0.f));
  }
  L->tid.resize(T);
  L->splatted.resize(T);
  L->mvT.assign(16, 0.f); L->projT.assign(16, 0.f); L->fxy.assign(2, 0.f);
  for (std::size_t i = 0; i < n; ++i) {
    const std::size_t t = std::min<std::size_t>(i / perTile, T - 1), s = i - t * perTile;
    std::memcpy(&L->vertsIn[t][s * S], g16 + i * 16, 16 * sizeof(float));
  }
  const TiledFramebuffer tfb(IPU_TILEWIDTH, IPU_TILEHEIGHT);
  for (int t = 0; t < T; ++t) {
    GSplat &v = L->verts[t];
    L->tid[t][0] = t; L->splatted[t][0] = 0;
    v.modelView.bind(L->mvT.data(), 16);
    v.projection.bind(L->projT.data(), 16);
    v.fxy.bind(L->fxy.data(), 2);
    v.tile_id.bind(L->tid[t].data(), 1);
    v.splatted.bind(L->splatted[t].data(), 1);
    v.vertsIn.bind(L->vertsIn[t].data(), L->vertsIn[t].size());
    v.indices.bind(L->indices[t].data(), L->indices[t].size());
    v.gaus2D.bind(L->gaus2D[t].data(), L->gaus2D[t].size());
    v.localFb.bind(L->fb[t].data(), L->fb[t].size());
    v.rightOut.bind(L->out[0][t].data(), L->out[0][t].size());
    v.leftOut.bind(L->out[1][t].data(), L->out[1][t].size());
    v.upOut.bind(L->out[2][t].data(), L->out[2][t].size());
    v.downOut.bind(L->out[3][t].data(), L->out[3][t].size());
    v.rightIn.bind(L->in[0][t].data(), L->in[0][t].size());
    v.leftIn.bind(L->in[1][t].data(), L->in[1][t].size());
    v.upIn.bind(L->in[2][t].data(), L->in[2][t].size());
    v.downIn.bind(L->in[3][t].data(), L->in[3][t].size());
    // edge_builder.cpp:47-81 (0=right 1=left 2=up 3=down)
    auto b = tfb.checkImageBoundaries(t);
    if (b.up) L->edges.push_back({t, 2, t, 2});
    if (b.left) L->edges.push_back({t, 1, t, 1});
    if (b.right) L->edges.push_back({t, 0, t, 0});
    if (b.down) L->edges.push_back({t, 3, t, 3});
    if (!b.right && t + 1 < T) {
      L->edges.push_back({t, 0, t + 1, 1});       // rightOut -> leftIn
      L->edges.push_back({t + 1, 1, t, 0});       // leftOut  -> rightIn
    } else if (!b.right) {
      L->edges.push_back({t, 0, t, 0});
    }
    if (!b.down && t + TA < T) {
      L->edges.push_back({t, 3, t + TA, 2});      // downOut -> upIn
      L->edges.push_back({t + TA, 2, t, 3});      // upOut   -> downIn
    } else if (!b.down) {
      L->edges.push_back({t, 3, t, 3});
    }
  }
  return L;
}

// One "project" program run (ipu_rasteriser.cpp:393-399): broadcast MVP, Execute(splatCs), copies.
void ref_lattice_step(void *h, const float *mv, const float *proj, float fov, float scale_div,
                      int nthreads) {
  auto *L = static_cast<Lattice *>(h);
  set_geometry(L->W, L->H, L->TA, L->TD);
  if (nthreads <= 0) nthreads = omp_get_max_threads();
  const glm::mat4 viewT = glm::transpose(glm::make_mat4(mv));     // ipu_rasteriser.cpp:86-102
  const glm::mat4 projT = glm::transpose(glm::make_mat4(proj));
  std::memcpy(L->mvT.data(), glm::value_ptr(viewT), 64);
  std::memcpy(L->projT.data(), glm::value_ptr(projT), 64);
  L->fxy[0] = fov; L->fxy[1] = scale_div;
  double t0 = now_ms();
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 1)
  for (int t = 0; t < L->numTiles; ++t) {
    // workers 1..5 only clear (codelets.cpp:608-617); run them first, then worker 0.
    for (unsigned w = 1; w < 6; ++w) L->verts[t].compute(w);
    L->verts[t].compute(0);
  }
  L->compute_ms = now_ms() - t0;
  t0 = now_ms();
  for (const auto &e : L->edges)                                   // edge_builder.cpp:32
    std::memcpy(L->in[e.dstDir][e.dst].data(), L->out[e.srcDir][e.src].data(),
                L->out[e.srcDir][e.src].size() * sizeof(float));
  L->exchange_ms = now_ms() - t0;
}

// Per-tile render lists of the last step (gid-1 of each gaus2D record is not stored in
// Gaussian2D, so lists are recovered by matching (mean,z) is impossible; instead we report the
// held Gaussian3D gids whose record was appended: re-derive with the same predicate the vertex
// uses, codelets.cpp:470,493).
void ref_lattice_get(void *h, uint32_t *counts, float *fb_tiled, double *ms2) {
  auto *L = static_cast<Lattice *>(h);
  for (int t = 0; t < L->numTiles; ++t) {
    counts[t] = L->splatted[t][0];
    std::memcpy(fb_tiled + std::size_t(t) * L->fb[t].size(), L->fb[t].data(), L->fb[t].size() * sizeof(float));
  }
  ms2[0] = L->compute_ms; ms2[1] = L->exchange_ms;
}

// Held gids per tile (vertsIn occupancy) after the last step: tile t holds gid iff a slot has it.
std::size_t ref_lattice_held(void *h, int t, int32_t *gids, std::size_t cap) {
  auto *L = static_cast<Lattice *>(h);
  const std::size_t S = sizeof(Gaussian3D);
  std::size_t k = 0;
  for (std::size_t s = 0; s < L->slotsPerTile; ++s) {
    float gid = L->vertsIn[t][s * S + 15];
    if (gid > 0) { if (k < cap) gids[k] = int32_t(gid) - 1; ++k; }
  }
  return k;
}
// Rendered list of tile t after the last step: the first splatted[0] records of gaus2D (sorted
// by renderTile's quicksort, codelets.cpp:360) as (mean.x, mean.y, z) triples.  Gaussian2D keeps
// no gid (ipu_geometry.hpp:232-236), so the caller matches entries by these values.
std::size_t ref_lattice_rendered(void *h, int t, float *xyz, std::size_t cap) {
  auto *L = static_cast<Lattice *>(h);
  const std::size_t S = sizeof(Gaussian2D);          // stride in FLOATS (codelets.cpp:370,494)
  const std::size_t n = L->splatted[t][0];
  for (std::size_t k = 0; k < n && k < cap; ++k) {
    Gaussian2D g;
    std::memcpy(&g, &L->gaus2D[t][k * S], sizeof(g));
    xyz[k * 3 + 0] = g.mean.x; xyz[k * 3 + 1] = g.mean.y; xyz[k * 3 + 2] = g.z;
  }
  return n;
}
void ref_lattice_destroy(void *h) { delete static_cast<Lattice *>(h); }

} // extern "C"
