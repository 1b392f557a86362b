// gsplat_api.cu — the C ABI of libgsplat_b200.so (include/gsplat_b200.h): context, memory, the
// per-frame stage sequence K1 -> sort A -> K2 -> sort B -> K4 -> K5, read-back.  Host side of what
// splat::IpuSplatter + ipu_utils::GraphManager do for the reference
// (src/splat/ipu_rasteriser.cpp:49-144,250-420; include/ipu/ipu_utils.hpp:523-623).
#include "../../include/gsplat_b200.h"
#include "gsplat_internal.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

using namespace gsplat;

namespace {
thread_local std::string g_create_error;

// stage events of a frame: start, K1, then {select + sort A, K2, sort B, K4, K5} per depth slice
constexpr int MAX_SLICES = GSPLAT_MAX_DEPTH_SLICES;
enum Ev { EV_START = 0, EV_K1, EV_SLICE0, EV_COUNT = EV_SLICE0 + 5 * MAX_SLICES };
enum SliceEv { SEV_SORTA = 0, SEV_K2, SEV_SORTB, SEV_K4, SEV_K5, SEV_COUNT };
}  // namespace

// One in-flight frame: its stage events and the pinned words its scalars are copied to.  Frames
// are enqueued without any host synchronisation; a slot is resolved (event wait, overflow check,
// timing fold) when it is reused, when a later call needs the frame's results, or opportunistically.
constexpr int NUM_SLOTS = 4;
struct FrameSlot {
  cudaEvent_t ev[EV_COUNT]{};
  cudaEvent_t done = nullptr;
  uint32_t *h = nullptr;   // pinned copy of the frame's device scalars (enum Scalar)
  int num_slices = 1;
  cudaEvent_t t_begin = nullptr, t_end = nullptr;   // whole-frame device time (slice auto-tuning)
  bool auto_timed = false;
  bool pending = false, timed = false, overflowed = false;
  uint64_t frame_no = 0;
};

// One slot of the double-buffered frame stream (gsplat_render_streamed / gsplat_acquire_frame):
// device staging written on the render stream, pinned host copies filled on the copy stream.
struct StreamSlot {
  uint8_t *d_bgr = nullptr, *h_bgr = nullptr;
  uint32_t *d_counts = nullptr, *h_counts = nullptr;
  cudaEvent_t rendered = nullptr, copied = nullptr;
  uint64_t ticket = 0, frame_no = 0;
  bool used = false;
  // verdict of the frame this slot carries, latched when its FrameSlot resolves: the FrameSlot ring
  // may be recycled (overflow re-runs consume extra slots) before the ticket is acquired
  bool resolved = false, overflowed = false;
  FrameParams fp{};
};

// device scalars
// device scalars; words [0, SC_FRAME_WORDS) are zeroed at the start of every frame and copied to
// the frame slot's pinned words at its end
enum Scalar { SC_V = 0, SC_OVERFLOW = 1, SC_BAND0 = 2 /* = SC_V + BAND_WORD_OFFSET: {row_begin, row_end}, written by K1 */,
              SC_TICKET0 = 8 /* K2 chain ticket per slice */,
              SC_DS0 = 16 /* duplicates emitted per slice */, SC_NS0 = 24 /* Gaussians selected per slice */,
              SC_SELTICKET0 = 32 /* selection chain ticket per slice */, SC_TAU0 = 40 /* depth-key thresholds */,
              SC_KRANGE0 = 48 /* {max, max of ~} depth key of each slice's selection (2 words per slice) */,
              SC_FRAME_WORDS = 64 };
static_assert(SC_BAND0 == SC_V + gsplat::BAND_WORD_OFFSET, "K1 writes the band next to its drawn counter");
enum Persist { P_N = 0 /* scene size */, P_SORT_N = 1 /* gsplat_sort_pairs count */, P_COUNT = 16 };

// A captured frame (CUDA graph): the whole stage sequence of one frame for one slicing mode.  Only
// K1's camera argument changes from frame to frame; it is refreshed in place before each launch.
struct FrameGraph {
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  cudaGraphNode_t k1 = nullptr;
  cudaKernelNodeParams k1_params{};
  int sorted_a = 0, sorted = 0;
  unsigned long long launches = 0;
};

struct gsplat_ctx {
  gsplat_config cfg{};
  cudaStream_t st = nullptr;
  cudaStream_t aux_st = nullptr;            // per-tile counts run beside the slices (nothing in the frame reads them)
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  std::string err;
  int tiles_across = 0, tiles_down = 0, num_tiles = 0, tile_bits = 0;
  // scene (immutable after upload; the lanes of a dispatcher share lane 0's arrays: scene_owned = false)
  SceneSoA scene{};
  bool scene_owned = true;
  bool is_lane = false;      // one of a dispatcher's lanes (several frames in flight)
  // scale divisor (bits) the scene's cached ComputeCov3D results were computed for; the lanes of a
  // dispatcher read lane 0's (cov_state points at the owner's word)
  uint32_t cov_div_bits = 0xffffffffu;
  const uint32_t *cov_state = nullptr;
  // per-Gaussian frame buffers
  Projected proj{};
  float *dbg_rec8 = nullptr;
  int32_t *dbg_rect4 = nullptr;
  uint32_t *d_scalars = nullptr;
  uint32_t *h_scalars = nullptr;   // pinned, NUM_SLOTS * 4 words
  // sort A: Gaussians by depth (capacity N); sort B: duplicates by tile (capacity grows)
  // Frames in flight (gsplat_config.concurrent_frames > 1): this context is then only a dispatcher
  // over `lanes` complete contexts (own stream, own buffers, own copy of the scene) that take
  // the frames round-robin; result reads go to the lane of the frame rendered last.
  std::vector<gsplat_ctx *> lanes;
  int next_lane = 0, last_lane = 0;
  // Slicing mode under a dispatcher.  A lane cannot tune itself: with other frames in flight its
  // frame times measure the neighbours' load.  The dispatcher decides for all lanes from frames it
  // runs ALONE: every DISPATCH_AUTO_PERIOD frames it drains the lanes and renders one frame in each
  // mode by itself (the caller's own frames, in order), then keeps the faster mode (3 % hysteresis).
  int forced_mode = -1;             // lane: -1 = decide locally, 0 = one slice, 1 = sliced
  int d_mode = 1, d_phase = 0;      // dispatcher: mode in use; measured frames still to run
  int d_retries = 0;                // measurements repeated because the frame overflowed
  uint64_t d_frames = 0;
  double d_ms[2] = {0.0, 0.0};
  cudaEvent_t d_t0 = nullptr, d_t1 = nullptr;
  SortBuffers sort_a{}, sort_b{};
  int sorted_a = 0, sorted = 0;    // which of sort_x.keys/vals holds the sorted result
  unsigned long long *lookback = nullptr;
  uint32_t n_blocks = 0, n_sel_blocks = 0;
  uint2 *ranges = nullptr;
  uint32_t *counts = nullptr;      // a11: reference list length of every tile (tile_count_kernel)
  int *count_diff = nullptr;       // difference grid of tile_count_kernel (kept zero between frames)
  uint32_t *tile_order = nullptr;  // K5 launch order of the tiles (longest list first), per slice
  uint32_t *live_sat = nullptr;    // summed-area table of the tiles still being blended (depth slices)
  // depth slices (select.cu): slice s = depth keys in [tau[s-1], tau[s]), tau = sampled quantiles
  int num_slices = 1;              // slices of a sliced frame (plan_slices)
  int frame_slices = 1;            // slices of the frame being / last enqueued
  // depth_slices == 0: sliced and one-slice frames give the same image, so the context measures
  // both (device time, smallest of the last four samples) and renders with the faster; every AUTO_PROBE-th frame
  // re-measures the other
  bool auto_slices = false;
  double ema_ms[2] = {0, 0};       // [0] one slice  [1] sliced: current estimate (min of the last 4 samples)
  float samples[2][4] = {{0}};
  uint64_t nsamples[2] = {0, 0};
  bool ema_valid[2] = {false, false};
  uint64_t auto_frames = 0, mode_frames[2] = {0, 0};
  int auto_mode = -1;               // the mode the auto-tuner currently renders with (-1: none yet)
  SliceFractions slice_fr{};       // cumulative fractions of the thresholds (num_slices - 1 of them)
  unsigned long long *sel_lookback = nullptr;   // selection chains, one per slice
  // Everything that is zeroed at the start of every frame sits in two blocks so that a frame starts
  // with two memsets instead of six (each one is a node on the frame's critical path):
  //   frame_block (create): device scalars [0, SC_FRAME_WORDS) | TileState arrays | ranges of every slice
  //   chain_block (upload): look-back words of K2 and of the selection, every slice
  uint8_t *frame_block = nullptr, *chain_block = nullptr;
  size_t frame_block_bytes = 0, chain_block_bytes = 0;
  uint32_t *d_persist = nullptr;   // device scalars that outlive a frame (enum Persist)
  uint2 *ranges_last = nullptr;    // ranges of the last slice of the last frame
  uint8_t *tile_state = nullptr;   // inside frame_block: TileState arrays
  TileState ts{};                  // per tile / per part: every pixel has stopped
  int blend_split = 1;             // K5 CTAs per semantic tile in the first depth slice
  int blend_split_later = 1;       // ... and in the later slices (fewer live tiles: narrower parts fill the SMs)
  bool blend_wide = false;         // K5 in its 53-register form (one frame at a time), see blend.cu
  bool k1_dense = true;            // K1 at 32 registers / 64 warps per SM (one frame at a time), see project.cu
  // frame graphs: [0] one slice, [1] sliced; dropped whenever buffers, scene or band change
  FrameGraph graphs[2];
  bool use_graphs = true;          // GSPLAT_GRAPHS=0 disables (eager launches)
  uint8_t *l2_scratch = nullptr;   // gsplat_flush_l2 (allocated on first use)
  float *tstate = nullptr;         // per pixel: transmittance carried between slices (-1 = stopped)
  unsigned long long *pair_counters = nullptr;   // non-null while a counting re-run is enqueued
  // parity tap: per-slice sorted lists of the last frame (filled only with cfg.debug_records)
  std::vector<std::vector<uint2>> log_ranges;
  std::vector<std::vector<uint32_t>> log_vals;
  float4 *fb = nullptr;
  uint8_t *bgr = nullptr;
  // state
  FrameParams fp{};          // set_camera
  FrameParams frame_fp{};    // camera of the last enqueued frame (re-used when a frame is re-run)
  bool have_scene = false, have_camera = false, have_frame = false;
  bool last_injected = false;
  gsplat_frame_stats stats{};
  FrameSlot slots[NUM_SLOTS];
  uint64_t frame_no = 0;      // frames enqueued so far; the last one lives in slot (frame_no-1) % NUM_SLOTS
  size_t need_grow = 0;       // largest D seen by an overflowed frame
  // frame numbers that overflowed the duplicate buffer (ring): the verdict outlives the FrameSlot,
  // which overflow re-runs may recycle before a ticket that names the frame is acquired
  uint64_t ovf_log[64] = {0};
  uint64_t ovf_logged = 0;
  cudaEvent_t t0 = nullptr, t1 = nullptr, post0 = nullptr, post1 = nullptr;
  bool post_pending = false;
  // frame stream (allocated on first use)
  cudaStream_t copy_st = nullptr;
  StreamSlot sslots[GSPLAT_STREAM_SLOTS];
  uint64_t next_ticket = 1;
  float last_ms[GSPLAT_NUM_TIMINGS] = {0};
  double acc_ms[GSPLAT_NUM_TIMINGS] = {0};
  uint64_t acc_frames = 0;
};

namespace {

int fail(gsplat_ctx *c, int code, const std::string &msg) {
  if (c) c->err = msg; else g_create_error = msg;
  return code;
}

#define CU(call)                                                                                  \
  do {                                                                                            \
    cudaError_t e_ = (call);                                                                      \
    if (e_ != cudaSuccess)                                                                        \
      return fail(ctx, GSPLAT_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));        \
  } while (0)

template <typename T> void dfree(T *&p) { if (p) { cudaFree(p); p = nullptr; } }
void free_scene(SceneSoA &sc) {
  dfree(sc.pos_c); dfree(sc.cov_a); dfree(sc.cov_b); dfree(sc.rgba); dfree(sc.quat); dfree(sc.scl);
}
uint32_t float_bits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

void drop_graphs(gsplat_ctx *ctx) {
  for (auto &g : ctx->graphs) {
    if (g.exec) cudaGraphExecDestroy(g.exec);
    if (g.graph) cudaGraphDestroy(g.graph);
    g = FrameGraph{};
  }
}

void free_sort(SortBuffers &s) {
  dfree(s.keys[0]); dfree(s.keys[1]); dfree(s.vals[0]); dfree(s.vals[1]); dfree(s.block_hist); dfree(s.digit_totals);
  dfree(s.tile_state); dfree(s.sweep_ctl);
  s.capacity = 0; s.max_tiles = 0;
}

int alloc_sort(gsplat_ctx *ctx, SortBuffers &s, size_t capacity) {
  drop_graphs(ctx);   // captured frames hold the old buffers' addresses
  if (capacity > 0xfffff000ull) return fail(ctx, GSPLAT_E_CAPACITY, "element count exceeds 32-bit indexing");
  cudaSetDevice(ctx->cfg.device);
  free_sort(s);
  capacity = (capacity + RS_TILE - 1) / RS_TILE * RS_TILE;
  if (capacity == 0) capacity = RS_TILE;
  s.max_tiles = (uint32_t)(capacity / RS_TILE);
  for (int i = 0; i < 2; ++i) {
    CU(cudaMalloc(&s.keys[i], capacity * sizeof(uint32_t)));
    CU(cudaMalloc(&s.vals[i], capacity * sizeof(uint32_t)));
  }
  CU(cudaMalloc(&s.block_hist, (size_t)256 * s.max_tiles * sizeof(uint32_t)));
  CU(cudaMalloc(&s.digit_totals, 256 * sizeof(uint32_t)));
  CU(cudaMemset(s.digit_totals, 0, 256 * sizeof(uint32_t)));
  // The sort passes (radix_sort.cu) are single sweeps: 33 launches per frame instead of the 57 of the
  // three-kernel passes.  How many CTAs a sweep gets is the per-context choice (profiles/ab_sort_r2.txt,
  // 6 M bench frame): a frame that runs alone wants the whole GPU (3 CTAs per SM: sort A + B 0.340 ms
  // against 0.394 with one per SM), but with several frames in flight the CTAs that wait in the
  // look-back hold slots other lanes' kernels could use — 444 CTAs: 1574 frames/s, 148: 1637 (three-
  // kernel passes: 1607).  GSPLAT_SORT_SWEEP=0 selects the three-kernel passes, GSPLAT_SORT_SWEEP_GRID
  // sets the CTA count.
  bool sweep = true;
  if (const char *e = std::getenv("GSPLAT_SORT_SWEEP")) sweep = std::atoi(e) != 0;
  s.sweep_grid = ctx->is_lane ? 148u : (uint32_t)RS_GRID;
  if (const char *e = std::getenv("GSPLAT_SORT_SWEEP_GRID")) s.sweep_grid = (uint32_t)std::max(1, std::atoi(e));
  if (sweep) {
    CU(cudaMalloc(&s.tile_state, (size_t)256 * s.max_tiles * sizeof(unsigned long long)));
    CU(cudaMemset(s.tile_state, 0, (size_t)256 * s.max_tiles * sizeof(unsigned long long)));
    CU(cudaMalloc(&s.sweep_ctl, RS_SWEEP_CTL_WORDS * sizeof(uint32_t)));
    CU(cudaMemset(s.sweep_ctl, 0, RS_SWEEP_CTL_WORDS * sizeof(uint32_t)));
  }
  s.capacity = (uint32_t)capacity;
  return GSPLAT_OK;
}

// One-time kernel setup per DEVICE (not per process): the dynamic-shared-memory opt-in is stored in
// the device's context, so a second context on another GPU of the same process needs its own.
// Thread-safe: contexts may be created from several host threads (one per GPU).
cudaError_t setup_device_kernels(int device) {
  static std::mutex mu;
  static bool done[256] = {false};
  std::lock_guard<std::mutex> lock(mu);
  if (device < 0 || device >= 256) return cudaErrorInvalidDevice;
  if (done[device]) return cudaSuccess;
  cudaError_t e = cudaSetDevice(device);
  if (e == cudaSuccess) e = setup_blend_kernels();
  if (e == cudaSuccess) e = setup_duplicate_kernels();
  if (e == cudaSuccess) e = setup_sort_kernels();
  if (e == cudaSuccess) done[device] = true;
  return e;
}

int ceil_log2(uint32_t v) { int b = 0; while ((1u << b) < v) ++b; return b; }

// Depth slices: cumulative fractions of the drawn Gaussians (nearest first) at which the frame is
// cut.  The first slice is small (dense tiles saturate within their first few thousand Gaussians:
// 8 % with the default two slices), further cuts are spaced geometrically up to 1.
// depth_slices == 1: one slice (every tile's list complete, the reference's lists).  Scenes too
// small to be worth cutting get one slice.  GSPLAT_SLICE_FRACTIONS="f0,f1,..." overrides
// (experiments; the last slice always ends at 1).
void plan_slices(gsplat_ctx *ctx) {
  const uint32_t n = ctx->scene.n;
  int want = ctx->cfg.depth_slices ? ctx->cfg.depth_slices : GSPLAT_DEFAULT_DEPTH_SLICES;
  if (!ctx->cfg.depth_slices && n < 65536) want = 1;
  double frac[MAX_SLICES];
  for (int s = 0; s < want; ++s) frac[s] = std::pow(0.08, (double)(want - 1 - s) / (want > 2 ? 0.5 * (want - 1) : 1.0));
  if (const char *e = std::getenv("GSPLAT_SLICE_FRACTIONS")) {
    if (!ctx->cfg.depth_slices) {
      int k = 0;
      const char *q = e;
      while (*q && k < MAX_SLICES - 1) {
        char *end = nullptr;
        const double v = std::strtod(q, &end);
        if (end == q) break;
        // cut points must increase: a non-monotone list would make the slices' key intervals
        // overlap and Gaussians would be binned and blended twice
        if (v > 0.0 && v < 1.0 && (k == 0 || v > frac[k - 1])) frac[k++] = v;
        q = *end == ',' ? end + 1 : end;
      }
      want = k + 1;
      frac[k] = 1.0;
    }
  }
  ctx->num_slices = want;
  ctx->frame_slices = want;
  ctx->auto_slices = !ctx->cfg.depth_slices && want > 1 && !std::getenv("GSPLAT_SLICE_FRACTIONS");
  ctx->ema_valid[0] = ctx->ema_valid[1] = false;
  ctx->nsamples[0] = ctx->nsamples[1] = 0;
  ctx->auto_frames = 0;
  ctx->auto_mode = -1;
  ctx->slice_fr.count = want - 1;
  for (int s = 0; s + 1 < want; ++s) ctx->slice_fr.f[s] = (float)frac[s];
}

// glm type_mat4x4.inl:630-648, left-to-right sums; column-major 16-float arrays
void mat4_mul(const float *a, const float *b, float *r) {
  for (int c = 0; c < 4; ++c)
    for (int k = 0; k < 4; ++k)
      r[c * 4 + k] = ((a[0 + k] * b[c * 4 + 0] + a[4 + k] * b[c * 4 + 1]) + a[8 + k] * b[c * 4 + 2]) +
                     a[12 + k] * b[c * 4 + 3];
}

}  // namespace

extern "C" {

// ---- frames in flight: routing of the C ABI over the lanes of a dispatcher context ---------------
// ROUTE_LAST: the call refers to the frame rendered last (reads, stats, parity taps).
// ROUTE_ALL:  the call sets state every lane needs (scene, camera, band) or waits for all of them.
#define ROUTE_LAST(call)                                              \
  if (ctx && !ctx->lanes.empty()) {                                   \
    gsplat_ctx *outer_ = ctx;                                         \
    ctx = outer_->lanes[outer_->last_lane];                           \
    const int rc_ = (call);                                           \
    if (rc_ < 0) outer_->err = ctx->err;                              \
    return rc_;                                                       \
  }
#define ROUTE_ALL(call)                                               \
  if (ctx && !ctx->lanes.empty()) {                                   \
    gsplat_ctx *outer_ = ctx;                                         \
    for (gsplat_ctx *lane_ : outer_->lanes) {                         \
      ctx = lane_;                                                    \
      const int rc_ = (call);                                         \
      if (rc_ < 0) { outer_->err = lane_->err; return rc_; }          \
    }                                                                 \
    return GSPLAT_OK;                                                 \
  }

int gsplat_abi_version(void) { return GSPLAT_ABI_VERSION; }

int gsplat_config_default(gsplat_config *cfg) {
  if (!cfg) return GSPLAT_E_ARG;
  std::memset(cfg, 0, sizeof *cfg);
  cfg->device = 0;
  cfg->width = 1280; cfg->height = 720;      // tile_config.hpp:5-6
  cfg->tile_w = 32; cfg->tile_h = 20;        // tile_config.hpp:9-15 (1280/40, 720/36)
  cfg->guard_factor = 15.0f;                 // codelets.cpp:622
  cfg->band_row_end = -1;                    // all tile rows
  return GSPLAT_OK;
}

const char *gsplat_last_error(const gsplat_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

void gsplat_destroy(gsplat_ctx *ctx) {
  if (!ctx) return;
  if (!ctx->lanes.empty()) {
    // lane 0 owns the scene arrays the others read: it goes last
    for (size_t l = ctx->lanes.size(); l-- > 0;) gsplat_destroy(ctx->lanes[l]);
    if (ctx->d_t0) cudaEventDestroy(ctx->d_t0);
    if (ctx->d_t1) cudaEventDestroy(ctx->d_t1);
    delete ctx;
    return;
  }
  if (std::getenv("GSPLAT_DEBUG"))
    std::fprintf(stderr, "[gsplat] auto=%d frames one-slice %llu (%.3f ms) sliced %llu (%.3f ms)\n", (int)ctx->auto_slices,
                 (unsigned long long)ctx->mode_frames[0], ctx->ema_ms[0], (unsigned long long)ctx->mode_frames[1], ctx->ema_ms[1]);
  cudaSetDevice(ctx->cfg.device);
  if (ctx->st) cudaStreamSynchronize(ctx->st);
  drop_graphs(ctx);
  if (ctx->scene_owned) free_scene(ctx->scene);
  dfree(ctx->proj.rec); dfree(ctx->proj.rect);
  dfree(ctx->dbg_rec8); dfree(ctx->dbg_rect4); dfree(ctx->frame_block); dfree(ctx->d_persist); dfree(ctx->chain_block);
  dfree(ctx->proj.depth_key);
  free_sort(ctx->sort_a); free_sort(ctx->sort_b);
  dfree(ctx->counts); dfree(ctx->fb); dfree(ctx->bgr);
  dfree(ctx->l2_scratch);
  dfree(ctx->count_diff); dfree(ctx->live_sat); dfree(ctx->tile_order); dfree(ctx->tstate);
  if (ctx->copy_st) cudaStreamSynchronize(ctx->copy_st);
  for (auto &ss : ctx->sslots) {
    dfree(ss.d_bgr); dfree(ss.d_counts);
    if (ss.h_bgr) cudaFreeHost(ss.h_bgr);
    if (ss.h_counts) cudaFreeHost(ss.h_counts);
    if (ss.rendered) cudaEventDestroy(ss.rendered);
    if (ss.copied) cudaEventDestroy(ss.copied);
  }
  if (ctx->copy_st) cudaStreamDestroy(ctx->copy_st);
  if (ctx->h_scalars) cudaFreeHost(ctx->h_scalars);
  for (auto &sl : ctx->slots) {
    for (auto &e : sl.ev) if (e) cudaEventDestroy(e);
    if (sl.done) cudaEventDestroy(sl.done);
    if (sl.t_begin) cudaEventDestroy(sl.t_begin);
    if (sl.t_end) cudaEventDestroy(sl.t_end);
  }
  if (ctx->t0) cudaEventDestroy(ctx->t0);
  if (ctx->t1) cudaEventDestroy(ctx->t1);
  if (ctx->post0) cudaEventDestroy(ctx->post0);
  if (ctx->post1) cudaEventDestroy(ctx->post1);
  if (ctx->aux_st) { cudaStreamSynchronize(ctx->aux_st); cudaStreamDestroy(ctx->aux_st); }
  if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
  if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
  if (ctx->st) cudaStreamDestroy(ctx->st);
  delete ctx;
}

int gsplat_create(const gsplat_config *cfg, gsplat_ctx **out) {
  gsplat_ctx *ctx = nullptr;
  if (!cfg || !out) return fail(nullptr, GSPLAT_E_ARG, "gsplat_create: null argument");
  *out = nullptr;
  if (cfg->width <= 0 || cfg->height <= 0 || cfg->tile_w <= 0 || cfg->tile_h <= 0 ||
      cfg->width % cfg->tile_w || cfg->height % cfg->tile_h)
    return fail(nullptr, GSPLAT_E_ARG, "gsplat_create: tile size must divide the image size");
  if (cfg->tile_w != 32 || cfg->tile_h % 4 || cfg->tile_h > 32)
    return fail(nullptr, GSPLAT_E_ARG, "gsplat_create: K5 supports tile_w == 32 and tile_h in {4,8,...,32}");
  if (cfg->width / cfg->tile_w > 65535 || cfg->height / cfg->tile_h > 65535)
    return fail(nullptr, GSPLAT_E_ARG, "gsplat_create: tile grid exceeds 16-bit tile coordinates");
  if (tile_count_smem(cfg->width / cfg->tile_w, cfg->height / cfg->tile_h) > TILE_COUNT_MAX_SMEM)
    return fail(nullptr, GSPLAT_E_ARG, "gsplat_create: tile grid too large for the per-tile count kernel");
  if (cfg->depth_slices < 0 || cfg->depth_slices > GSPLAT_MAX_DEPTH_SLICES)
    return fail(nullptr, GSPLAT_E_ARG, "gsplat_create: depth_slices must be 0 (default) .. GSPLAT_MAX_DEPTH_SLICES");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(nullptr, GSPLAT_E_CUDA, std::string("gsplat_create: no CUDA device (there is no CPU fallback): ") +
                                            cudaGetErrorString(e));
  if (cfg->device < 0 || cfg->device >= ndev) return fail(nullptr, GSPLAT_E_ARG, "gsplat_create: bad device ordinal");
  if (cfg->concurrent_frames < 0 || cfg->concurrent_frames > GSPLAT_MAX_CONCURRENT_FRAMES)
    return fail(nullptr, GSPLAT_E_ARG, "gsplat_create: concurrent_frames must be 0 (default) .. GSPLAT_MAX_CONCURRENT_FRAMES");
  {
    int nl = cfg->concurrent_frames;
    if (nl == 0) {
      if (cfg->collect_timings || cfg->debug_records) nl = 1;   // stage events and parity taps describe ONE frame at a time
      else {
        nl = GSPLAT_DEFAULT_CONCURRENT_FRAMES;
        if (const char *e2 = std::getenv("GSPLAT_CONCURRENT_FRAMES")) nl = std::atoi(e2);   // experiments
        if (nl < 1 || nl > GSPLAT_MAX_CONCURRENT_FRAMES) nl = 1;
      }
    }
    if (nl > 1) {
      gsplat_ctx *outer = new gsplat_ctx;
      outer->cfg = *cfg;
      gsplat_config lane_cfg = *cfg;
      lane_cfg.concurrent_frames = 1;
      for (int l = 0; l < nl; ++l) {
        gsplat_ctx *lane = nullptr;
        const int rc = gsplat_create(&lane_cfg, &lane);
        if (rc) {   // g_create_error is the lane's
          for (gsplat_ctx *made : outer->lanes) gsplat_destroy(made);
          delete outer;
          return rc;
        }
        lane->is_lane = true;
        if (!std::getenv("GSPLAT_K5_WIDE")) lane->blend_wide = false;   // several frames in flight: residency counts (blend.cu)
        if (!std::getenv("GSPLAT_K1_DENSE")) lane->k1_dense = false;    // ... and K1 must leave room beside it (project.cu)
        outer->lanes.push_back(lane);
      }
      cudaSetDevice(cfg->device);
      if (cudaEventCreate(&outer->d_t0) != cudaSuccess || cudaEventCreate(&outer->d_t1) != cudaSuccess) {
        gsplat_destroy(outer);
        return fail(nullptr, GSPLAT_E_CUDA, "gsplat_create: cudaEventCreate failed");
      }
      *out = outer;
      return GSPLAT_OK;
    }
  }
  ctx = new gsplat_ctx;
  ctx->cfg = *cfg;
  ctx->tiles_across = cfg->width / cfg->tile_w;
  ctx->tiles_down = cfg->height / cfg->tile_h;
  ctx->num_tiles = ctx->tiles_across * ctx->tiles_down;
  if (ctx->cfg.band_row_end < 0) ctx->cfg.band_row_end = ctx->tiles_down;   // -1: all rows ((0,0) is an EMPTY band)
  if (ctx->cfg.band_row_begin < 0 || ctx->cfg.band_row_end > ctx->tiles_down ||
      ctx->cfg.band_row_begin > ctx->cfg.band_row_end) {
    delete ctx;
    return fail(nullptr, GSPLAT_E_ARG, "gsplat_create: bad tile-row band");
  }
#define CUC(call)                                                                                  \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      std::string m_ = std::string(#call) + ": " + cudaGetErrorString(e_);                         \
      gsplat_destroy(ctx);                                                                         \
      return fail(nullptr, GSPLAT_E_CUDA, m_);                                                     \
    }                                                                                              \
  } while (0)
  CUC(cudaSetDevice(cfg->device));
  CUC(setup_device_kernels(cfg->device));
  CUC(cudaStreamCreateWithFlags(&ctx->st, cudaStreamNonBlocking));
  CUC(cudaStreamCreateWithFlags(&ctx->aux_st, cudaStreamNonBlocking));
  CUC(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
  CUC(cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming));
  CUC(cudaHostAlloc(&ctx->h_scalars, NUM_SLOTS * SC_FRAME_WORDS * sizeof(uint32_t), cudaHostAllocDefault));
  for (int k = 0; k < NUM_SLOTS; ++k) {
    FrameSlot &sl = ctx->slots[k];
    for (auto &ev : sl.ev) CUC(cudaEventCreate(&ev));
    CUC(cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
    CUC(cudaEventCreate(&sl.t_begin));
    CUC(cudaEventCreate(&sl.t_end));
    sl.h = ctx->h_scalars + SC_FRAME_WORDS * k;
  }
  CUC(cudaEventCreate(&ctx->t0));
  CUC(cudaEventCreate(&ctx->t1));
  CUC(cudaEventCreate(&ctx->post0));
  CUC(cudaEventCreate(&ctx->post1));
  {
    const size_t tsb = (tile_state_bytes(ctx->num_tiles) + 15) / 16 * 16;
    ctx->frame_block_bytes = SC_FRAME_WORDS * sizeof(uint32_t) + tsb + (size_t)MAX_SLICES * ctx->num_tiles * sizeof(uint2);
    CUC(cudaMalloc(&ctx->frame_block, ctx->frame_block_bytes));
    CUC(cudaMemset(ctx->frame_block, 0, ctx->frame_block_bytes));
    ctx->d_scalars = reinterpret_cast<uint32_t *>(ctx->frame_block);
    ctx->tile_state = ctx->frame_block + SC_FRAME_WORDS * sizeof(uint32_t);
    ctx->ranges = reinterpret_cast<uint2 *>(ctx->tile_state + tsb);
    ctx->ranges_last = ctx->ranges;
    CUC(cudaMalloc(&ctx->d_persist, P_COUNT * sizeof(uint32_t)));
    CUC(cudaMemset(ctx->d_persist, 0, P_COUNT * sizeof(uint32_t)));
  }
  ctx->tile_bits = ceil_log2((uint32_t)ctx->num_tiles);
  if (ctx->tile_bits < 1) ctx->tile_bits = 1;
  const size_t px = (size_t)cfg->width * cfg->height;
  CUC(cudaMalloc(&ctx->counts, sizeof(uint32_t) * ctx->num_tiles));
  CUC(cudaMalloc(&ctx->fb, sizeof(float4) * px));
  CUC(cudaMalloc(&ctx->bgr, px * 3));
  CUC(cudaMemset(ctx->fb, 0, sizeof(float4) * px));
  CUC(cudaMemset(ctx->bgr, 0, px * 3));
  CUC(cudaMemset(ctx->counts, 0, sizeof(uint32_t) * ctx->num_tiles));
  {
    const size_t grid = (size_t)(ctx->tiles_across + 1) * (ctx->tiles_down + 1);
    CUC(cudaMalloc(&ctx->live_sat, sizeof(uint32_t) * grid));
    CUC(cudaMalloc(&ctx->tile_order, sizeof(uint32_t) * ctx->num_tiles));
    CUC(cudaMalloc(&ctx->count_diff, sizeof(int) * grid));
    CUC(cudaMemset(ctx->count_diff, 0, sizeof(int) * grid));
    const size_t T4 = ((size_t)ctx->num_tiles + 3) / 4 * 4;
    ctx->ts.parts_done = reinterpret_cast<uint32_t *>(ctx->tile_state);
    ctx->ts.tile_done = ctx->tile_state + 4 * T4;
    ctx->ts.part_done = ctx->ts.tile_done + T4;
    // K5 CTAs per tile: two 16-column strips (GSPLAT_K5_SPLIT = 1, 2 or 4 overrides)
    int split = 2;
    if (const char *e = std::getenv("GSPLAT_K5_SPLIT")) split = std::atoi(e);
    if (split != 1 && split != 2 && split != 4) split = 2;
    ctx->blend_split = split;
    // later slices: also two strips (measured with this round's K5, profiles/ab_k5_r2e.txt: 2 strips 1603,
    // 4 strips 1582, 1 strip 1592 frames/s; the earlier K5 preferred 4)
    int later = 2;
    if (const char *e = std::getenv("GSPLAT_K5_SPLIT_LATER")) later = std::atoi(e);
    if (later != 1 && later != 2 && later != 4) later = 2;
    ctx->blend_split_later = later;
    // K5's 53-register form (blend.cu): for a frame that runs alone, and only while the launch is a few
    // waves of CTAs — its end, a handful of long-running CTAs, is then what counts (6 M @1080p: 0.311 ->
    // 0.292 ms; 1 M: 0.186 -> 0.171).  At 4K (26 k CTAs) the launch is throughput-bound and the 40-register
    // form with 4 CTAs per SM wins (0.66 against 0.69 ms).  Lanes of a dispatcher: switched off when the
    // lane is adopted (gsplat_create).
    ctx->blend_wide = (size_t)ctx->num_tiles * (size_t)split <= (size_t)148 * 64;
    if (const char *e = std::getenv("GSPLAT_K5_WIDE")) ctx->blend_wide = std::atoi(e) != 0;
    if (const char *e = std::getenv("GSPLAT_K1_DENSE")) ctx->k1_dense = std::atoi(e) != 0;
    if (const char *e = std::getenv("GSPLAT_GRAPHS")) ctx->use_graphs = std::atoi(e) != 0;
    CUC(cudaMalloc(&ctx->tstate, sizeof(float) * px));
  }
#undef CUC
  *out = ctx;
  return GSPLAT_OK;
}

// The scene's cached ComputeCov3D results (SceneSoA) belong to ONE scale divisor.  Called when the
// scene or the camera parameters of the context that OWNS the arrays change: recomputes them if the
// divisor differs, synchronously (rare: an upload, or the UI's scale slider).  Lanes that share the
// arrays are drained by the dispatcher before it routes the new camera (gsplat_set_camera).
static int refresh_cov3(gsplat_ctx *ctx) {
  if (!ctx->scene_owned || !ctx->have_scene || !ctx->have_camera) return GSPLAT_OK;
  const uint32_t want = float_bits(ctx->fp.scale_div);
  if (ctx->cov_div_bits == want) return GSPLAT_OK;
  CU(cudaSetDevice(ctx->cfg.device));
  CU(cudaStreamSynchronize(ctx->st));
  launch_cov3(ctx->scene, ctx->fp.scale_div, ctx->st);
  CU(cudaGetLastError());
  CU(cudaStreamSynchronize(ctx->st));
  ctx->cov_div_bits = want;
  return GSPLAT_OK;
}

// Scene upload of one (lane) context.  `shared` != nullptr: adopt another lane's device arrays
// instead of uploading a copy (the scene is read-only after upload; only the per-frame working
// buffers are per lane).
static int upload_scene_impl(gsplat_ctx *ctx, const float *g16, size_t n, gsplat_ctx *shared) {
  if (!ctx || (!g16 && n && !shared)) return fail(ctx, GSPLAT_E_ARG, "gsplat_upload_scene: null argument");
  if (n > 0x7fffffffull) return fail(ctx, GSPLAT_E_ARG, "gsplat_upload_scene: too many Gaussians");
  CU(cudaSetDevice(ctx->cfg.device));
  CU(cudaStreamSynchronize(ctx->st));
  drop_graphs(ctx);
  for (auto &sl : ctx->slots) sl.pending = false;
  if (ctx->scene_owned) free_scene(ctx->scene);
  ctx->scene = SceneSoA{};
  ctx->scene_owned = true;
  ctx->cov_div_bits = 0xffffffffu;
  ctx->cov_state = &ctx->cov_div_bits;
  dfree(ctx->proj.rec); dfree(ctx->proj.rect); dfree(ctx->chain_block);
  dfree(ctx->proj.depth_key);
  dfree(ctx->dbg_rec8); dfree(ctx->dbg_rect4);
  ctx->have_scene = ctx->have_frame = false;
  const size_t na = n ? n : 1;
  if (shared) {
    ctx->scene = shared->scene;
    ctx->scene_owned = false;
    ctx->cov_state = &shared->cov_div_bits;
  } else {
    ctx->scene.n = (uint32_t)n;
    CU(cudaMalloc(&ctx->scene.pos_c, na * sizeof(float4)));
    CU(cudaMalloc(&ctx->scene.cov_a, na * sizeof(float4)));
    CU(cudaMalloc(&ctx->scene.cov_b, na * sizeof(float4)));
    CU(cudaMalloc(&ctx->scene.rgba, na * sizeof(float4)));
    CU(cudaMalloc(&ctx->scene.quat, na * sizeof(float4)));
    CU(cudaMalloc(&ctx->scene.scl, na * sizeof(float4)));
  }
  CU(cudaMalloc(&ctx->proj.rec, na * 3 * sizeof(float4)));
  CU(cudaMemsetAsync(ctx->proj.rec, 0, na * 3 * sizeof(float4), ctx->st));
  CU(cudaMalloc(&ctx->proj.rect, na * sizeof(uint2)));
  ctx->n_blocks = (uint32_t)((n + K2_CHUNK - 1) / K2_CHUNK);
  ctx->n_sel_blocks = (uint32_t)((n + 4095) / 4096);
  ctx->chain_block_bytes = (size_t)MAX_SLICES * ((size_t)ctx->n_blocks + ctx->n_sel_blocks + 2) * sizeof(unsigned long long);
  CU(cudaMalloc(&ctx->chain_block, ctx->chain_block_bytes));
  ctx->lookback = reinterpret_cast<unsigned long long *>(ctx->chain_block);
  ctx->sel_lookback = ctx->lookback + (size_t)MAX_SLICES * (ctx->n_blocks + 1);
  CU(cudaMalloc(&ctx->proj.depth_key, na * sizeof(uint32_t)));
  plan_slices(ctx);
  if (ctx->cfg.debug_records) {
    CU(cudaMalloc(&ctx->dbg_rec8, na * 8 * sizeof(float)));
    CU(cudaMalloc(&ctx->dbg_rect4, na * 4 * sizeof(int32_t)));
  }
  {  // sort A holds the (depth key, index) pairs of the Gaussians selected for a slice: up to n
    int rc = alloc_sort(ctx, ctx->sort_a, n);
    if (rc) return rc;
    const uint32_t n32 = (uint32_t)n;
    CU(cudaMemcpyAsync(ctx->d_persist + P_N, &n32, 4, cudaMemcpyHostToDevice, ctx->st));
    CU(cudaStreamSynchronize(ctx->st));
  }
  if (n && !shared) {
    float *tmp = nullptr;
    CU(cudaMalloc(&tmp, n * 16 * sizeof(float)));
    cudaError_t e = cudaMemcpyAsync(tmp, g16, n * 16 * sizeof(float), cudaMemcpyHostToDevice, ctx->st);
    if (e == cudaSuccess) {
      launch_aos_to_soa(tmp, ctx->scene, ctx->st);
      e = cudaStreamSynchronize(ctx->st);
    }
    cudaFree(tmp);
    if (e != cudaSuccess) return fail(ctx, GSPLAT_E_CUDA, std::string("scene upload: ") + cudaGetErrorString(e));
  }
  if (ctx->sort_b.capacity == 0) {
    size_t cap = ctx->cfg.dup_capacity ? ctx->cfg.dup_capacity : (n * 4 + (size_t)RS_TILE);
    int rc = alloc_sort(ctx, ctx->sort_b, cap);
    if (rc) return rc;
  }
  ctx->have_scene = true;
  return refresh_cov3(ctx);
}

int gsplat_upload_scene(gsplat_ctx *ctx, const float *g16, size_t n) {
  if (ctx && !ctx->lanes.empty()) {   // one upload; the other lanes adopt lane 0's read-only arrays
    gsplat_ctx *first = ctx->lanes[0];
    for (size_t l = 1; l < ctx->lanes.size(); ++l) {   // they may hold pointers into the arrays about to be freed
      gsplat_ctx *lane = ctx->lanes[l];
      cudaSetDevice(lane->cfg.device);
      cudaStreamSynchronize(lane->st);
      if (!lane->scene_owned) { lane->scene = SceneSoA{}; lane->scene_owned = true; lane->have_scene = false; lane->cov_state = &lane->cov_div_bits; }
    }
    int rc = upload_scene_impl(first, g16, n, nullptr);
    if (rc < 0) { ctx->err = first->err; return rc; }
    for (size_t l = 1; l < ctx->lanes.size(); ++l) {
      rc = upload_scene_impl(ctx->lanes[l], nullptr, n, first);
      if (rc < 0) { ctx->err = ctx->lanes[l]->err; return rc; }
    }
    return GSPLAT_OK;
  }
  return upload_scene_impl(ctx, g16, n, nullptr);
}

int gsplat_set_camera(gsplat_ctx *ctx, const float mv[16], const float proj[16], float fov, float scale_div) {
  if (ctx && !ctx->lanes.empty() && ctx->lanes[0]->cov_div_bits != float_bits(scale_div)) {
    // the lanes share lane 0's covariance cache, which is about to be recomputed: drain them all
    for (gsplat_ctx *lane : ctx->lanes) { cudaSetDevice(lane->cfg.device); cudaStreamSynchronize(lane->st); }
  }
  ROUTE_ALL(gsplat_set_camera(ctx, mv, proj, fov, scale_div));
  if (!ctx || !mv || !proj) return fail(ctx, GSPLAT_E_ARG, "gsplat_set_camera: null argument");
  FrameParams &fp = ctx->fp;
  // codelets.cpp:443: mvp = projmatrix * viewmatrix (the host-side transpose of
  // ipu_rasteriser.cpp:86-102 and the codelet's transpose back, :625-628, cancel)
  mat4_mul(proj, mv, fp.mvp);
  // codelets.cpp:444-448: tan(0.5*fxy[0]) is a double expression narrowed to float; the focal
  // lengths use glm::tan on floats; tfb.width/height are uint16 (tile_config.hpp:22-31)
  fp.tanfov = (float)std::tan(0.5 * (double)fov);
  fp.focal_x = (float)(uint16_t)ctx->cfg.width / (2.f * std::tan(fov / 2.f));
  fp.focal_y = (float)(uint16_t)ctx->cfg.height / (2.f * std::tan(fov / 2.f));
  fp.scale_div = scale_div;
  fp.img_w = (float)ctx->cfg.width; fp.img_h = (float)ctx->cfg.height;
  fp.tile_w = (float)ctx->cfg.tile_w; fp.tile_h = (float)ctx->cfg.tile_h;
  // tb.diagonal().length() * clipSize (codelets.cpp:470,622; ipu_geometry.hpp:53-55,116-118)
  fp.guard = std::sqrt(fp.tile_w * fp.tile_w + fp.tile_h * fp.tile_h) * ctx->cfg.guard_factor;
  {
    // K1 tests the squared diagonal against guard_sq instead of taking a square root per Gaussian.
    // sqrtf is correctly rounded and monotone, so the set {q : sqrtf(q) >= guard} is an upper set of
    // the floats; its least element is found by stepping from guard * guard.
    float q = fp.guard * fp.guard;
    if (fp.guard > 0.f && std::isfinite(q)) {
      while (q > 0.f && std::sqrt(std::nextafter(q, 0.f)) >= fp.guard) q = std::nextafter(q, 0.f);
      while (std::isfinite(q) && std::sqrt(q) < fp.guard) q = std::nextafter(q, INFINITY);
    }
    if (!(fp.guard > 0.f)) q = 0.f;      // sqrtf(x) < guard never holds
    fp.guard_sq = q;
    int e = 0;
    fp.inv_tile_w = std::frexp(fp.tile_w, &e) == 0.5f ? 1.0f / fp.tile_w : 0.0f;
  }
  fp.tiles_across = ctx->tiles_across; fp.tiles_down = ctx->tiles_down;
  fp.row_begin = ctx->cfg.band_row_begin; fp.row_end = ctx->cfg.band_row_end;
  ctx->have_camera = true;
  return refresh_cov3(ctx);
}

int gsplat_set_band(gsplat_ctx *ctx, int row_begin, int row_end) {
  ROUTE_ALL(gsplat_set_band(ctx, row_begin, row_end));
  if (!ctx) return GSPLAT_E_ARG;
  if (row_begin < 0 || row_end > ctx->tiles_down || row_begin > row_end)
    return fail(ctx, GSPLAT_E_ARG, "gsplat_set_band: bad tile-row band");
  // no re-capture: the band reaches the frame through K1's argument block (refreshed per frame) and
  // from there, on the device, the kernels behind it
  const bool changed = row_begin != ctx->cfg.band_row_begin || row_end != ctx->cfg.band_row_end;
  ctx->cfg.band_row_begin = row_begin;
  ctx->cfg.band_row_end = row_end;
  ctx->fp.row_begin = row_begin;
  ctx->fp.row_end = row_end;
  if (changed) {
    // tiles that leave the band must read count 0 until the next frame's tile_count_finish_kernel
    // rewrites every count (rectangles are clamped to the band in K1, so outside tiles count 0)
    CU(cudaSetDevice(ctx->cfg.device));
    CU(cudaMemsetAsync(ctx->counts, 0, sizeof(uint32_t) * ctx->num_tiles, ctx->st));
  }
  return GSPLAT_OK;
}

}  // extern "C"

#include "gsplat_api_frame.inc"
#include "gsplat_group.inc"
