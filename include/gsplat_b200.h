/* gsplat_b200.h — C ABI of libgsplat_b200.so, the B200 (sm_100a) Gaussian-splat render core.
 *
 * This library replaces ONE path of Nmjfry/gaussian_splat_ipu: the per-frame project -> route to
 * tiles -> z-order -> alpha-blend work that the reference spreads over IPU tiles
 * (codelets/splat/codelets.cpp:143-641 `GSplat`, include/tileMapping/tile_config.hpp,
 * src/splat/ipu_rasteriser.cpp, src/splat/edge_builder.cpp).  The reference has no FFI layer; its
 * boundary for this path is the C++ class `splat::IpuSplatter`
 * (include/splat/ipu_rasteriser.hpp:20-55).  Each entry point below names the IpuSplatter member
 * (or codelet field) it stands in for.  All paths are relative to the reference tree.
 *
 * Conventions
 *   - plain C, no exceptions: every call returns GSPLAT_OK (0) or a negative gsplat_status;
 *     gsplat_last_error() gives the message of the last failure on that context
 *     (reference: std::runtime_error / poplar exceptions, include/ipu/ipu_utils.hpp:531-546).
 *   - a context owns one CUDA device, one stream and all device / pinned memory; every pointer
 *     argument is caller-owned and only touched during the call.
 *   - calls on one context are not re-entrant; distinct contexts are independent (one per GPU /
 *     view).  gsplat_render() is asynchronous on the context's stream; gsplat_read_*() and
 *     gsplat_sync() wait for it.
 *   - matrices are 16 floats in glm memory order (column-major, what glm::value_ptr yields).
 *   - there is NO CPU fallback: without a CUDA device gsplat_create() fails with GSPLAT_E_CUDA.
 */
#ifndef GSPLAT_B200_H
#define GSPLAT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSPLAT_ABI_VERSION 4

typedef enum gsplat_status {
  GSPLAT_OK = 0,
  GSPLAT_E_ARG = -1,      /* bad argument / call order                                   */
  GSPLAT_E_CUDA = -2,     /* CUDA runtime error (message has the cudaError string)       */
  GSPLAT_E_CAPACITY = -3, /* duplicate buffer could not be grown (never a silent drop;   */
                          /* the reference drops silently, codelets.cpp:565-569)         */
  GSPLAT_E_STATE = -4,    /* e.g. render before upload_scene / set_camera                */
  GSPLAT_E_NCCL = -5      /* NCCL unavailable or failed (message has ncclGetErrorString)  */
} gsplat_status;

typedef struct gsplat_ctx gsplat_ctx;

/* Geometry and constants of the path.  Defaults = the reference's compile-time values. */
typedef struct gsplat_config {
  int device;            /* CUDA ordinal                                                       */
  int width, height;     /* image size; IMWIDTH/IMHEIGHT, tile_config.hpp:5-6 (1280x720)      */
  int tile_w, tile_h;    /* semantic tile; IPU_TILEWIDTH/HEIGHT, tile_config.hpp:9-15 (32x20); */
                         /* must divide width/height                                          */
  float guard_factor;    /* clipSize, codelets.cpp:622 (15): bbox diagonal < 15 tile diagonals */
  int band_row_begin;    /* tile-row band [begin,end) this context renders (multi-GPU tile    */
  int band_row_end;      /* bands); end = -1 (the default) = all rows; begin ==   */
                         /* end is an empty band (a rank that got no rows)                    */
  int collect_timings;   /* record per-stage cudaEvents (gsplat_get_timings)                  */
  int debug_records;     /* K1 also writes the projected records for gsplat_read_projected    */
  size_t dup_capacity;   /* initial capacity of the (tile,Gaussian) duplicate buffers;        */
                         /* 0 = choose from the scene size; grows on demand                   */
  int depth_slices;      /* the frame is sorted, binned and blended in this many slices of     */
                         /* increasing depth; tiles whose pixels have all stopped (test_T <    */
                         /* 1e-4, codelets.cpp:405-408) drop out of the later slices, and so do */
                         /* the Gaussians that only cover such tiles.                           */
                         /* 1 = one slice: every tile's list is the reference's complete list.  */
                         /* 0 = default (GSPLAT_DEFAULT_DEPTH_SLICES).  Framebuffer and counts   */
                         /* do not depend on it; lists are exact prefixes of the reference's.   */
  int concurrent_frames; /* frames the context keeps in flight on the device (1..8): the frame is a */
                         /* chain of short, latency-bound launches, so consecutive frames run on    */
                         /* separate streams with their own working buffers ("lanes", round-robin)  */
                         /* and interleave on the SMs.  Every call that reads a result refers to    */
                         /* the frame rendered last; the depth_slices = 0 choice is then made for   */
                         /* all lanes from frames run alone.  0 = default: GSPLAT_DEFAULT_CONCURRENT_FRAMES, */
                         /* or 1 when collect_timings / debug_records are set.  Costs one set of    */
                         /* buffers (about 250 bytes per Gaussian) per frame in flight.  A context  */
                         /* with ONE frame in flight sorts with single-sweep passes (33 launches a  */
                         /* frame instead of 57: lower latency).                */
} gsplat_config;
#define GSPLAT_MAX_CONCURRENT_FRAMES 8
#define GSPLAT_DEFAULT_CONCURRENT_FRAMES 4
#define GSPLAT_MAX_DEPTH_SLICES 8
#define GSPLAT_DEFAULT_DEPTH_SLICES 2

/* Fills *cfg with the reference defaults (1280x720, 32x20, guard 15, device 0). */
int gsplat_config_default(gsplat_config *cfg);

/* IpuSplatter::IpuSplatter + GraphManager::compileOrLoad/prepareEngine
 * (ipu_rasteriser.cpp:49-83,250-406; ipu_utils.hpp:548-610): device setup, buffers. */
int gsplat_create(const gsplat_config *cfg, gsplat_ctx **out);
void gsplat_destroy(gsplat_ctx *ctx);
const char *gsplat_last_error(const gsplat_ctx *ctx); /* ctx may be NULL: last create() error */

/* Scene upload == the "write_verts" program (ipu_rasteriser.cpp:59-68,401-405,409-418).
 * `gaussians16` holds n records of the reference's Gaussian3D (ipu_geometry.hpp:305-311):
 * mean(x,y,z,w) colour(r,g,b,opacity) rot(w,x,y,z) scale(3, log) gid.  Records are converted to
 * the SoA layout on the device; gid <= 0 marks an empty slot (codelets.cpp:456), any gid > 0 is
 * treated as index+1 (src/main/splat.cpp:161). */
int gsplat_upload_scene(gsplat_ctx *ctx, const float *gaussians16, size_t n);

/* updateModelView + updateProjection + updateFocalLengths (ipu_rasteriser.cpp:86-110): the
 * per-frame inputs.  fov = fxy[0] (radians), scale_div = fxy[1] (src/main/splat.cpp:262). */
int gsplat_set_camera(gsplat_ctx *ctx, const float modelview[16], const float projection[16],
                      float fov, float scale_div);

/* Multi-GPU tile bands (the reference is single-IPU, src/main/splat.cpp:41-46): restrict
 * this context to the tile rows [row_begin, row_end) from the next gsplat_render() on.  K1 still
 * projects every Gaussian but clamps rectangles to the band, so K2-K5 only process the band's
 * share of the duplicates; tiles outside the band keep count 0 and their framebuffer chunks are
 * not written.  Band boundaries are host-side ints, re-chosen per frame from the previous frame's
 * tile counts by the caller. */
int gsplat_set_band(gsplat_ctx *ctx, int row_begin, int row_end);

/* One run of the "project" program (ipu_rasteriser.cpp:393-399,408-420), converged
 * (SURVEY §8a): K1 project/cull, K3a radix sort of the Gaussians by depth, K2 duplicate (in depth
 * order), K3b stable radix sort of the duplicates by tile, K4 tile ranges, K5 blend.
 * Enqueues the whole frame without waiting for the GPU.  If the frame turns out to need more
 * duplicate slots than allocated, the next gsplat_sync()/gsplat_read_*() grows the buffers and
 * re-runs it (same camera) before returning, so results that are read are always complete. */
int gsplat_render(gsplat_ctx *ctx);
int gsplat_sync(gsplat_ctx *ctx);

/* outputFramebuffer stream (ipu_rasteriser.cpp:280-285,398): width*height*4 floats, tile-major:
 * tile t = row*tiles_across + col holds tile_w*tile_h RGBA pixels, row-major inside the tile. */
int gsplat_read_framebuffer_tiled_f32(gsplat_ctx *ctx, float *dst);
/* IpuSplatter::getFrameBuffer (ipu_rasteriser.cpp:115-144): min(v*255,255) -> u8 (round half
 * even, clamp >= 0) -> RGBA2BGR -> de-tiled rows; dst is height rows of `pitch` bytes (>= 3*width).
 * The conversion runs on the device; only the BGR8 image crosses PCIe. */
int gsplat_read_framebuffer_bgr8(gsplat_ctx *ctx, uint8_t *dst, size_t pitch);
/* counts stream / IpuSplatter::getIPUHistogram (ipu_rasteriser.cpp:104-106,321-324,399;
 * codelets.cpp:501-504): Gaussians in each tile's render list, tiles_across*tiles_down values
 * (tiles outside the band are 0). */
int gsplat_read_tile_counts(gsplat_ctx *ctx, uint32_t *dst);

/* ---- double-buffered frame stream ---------------------------------------------------------------
 * The render loop of the reference renders into one cv::Mat while its AsyncTask thread encodes /
 * sends the previous one, then swaps them (src/main/splat.cpp:215-226,281-283: imagePtr /
 * imagePtrBuffered).  The same pipeline, GSPLAT_STREAM_SLOTS deep, with the device-to-host copy inside it:
 *
 *   gsplat_render_streamed   gsplat_render() + the BGR8 post-process, then an asynchronous copy of
 *                            the BGR8 image and the per-tile counts into one of the context's
 *                            GSPLAT_STREAM_SLOTS pinned host slots on a second (copy) stream.
 *                            Returns at once with a ticket; frame k+1 renders while frame k's
 *                            pixels cross PCIe.
 *   gsplat_acquire_frame     waits for that ticket's copy and hands out pointers INTO the pinned
 *                            slot: *bgr8 = height rows of 3*width bytes (getFrameBuffer's image),
 *                            *counts = tiles_across*tiles_down values (getIPUHistogram).  Either may
 *                            be NULL.  The pointers stay valid until the third-next
 *                            gsplat_render_streamed (the slot is then re-used); a ticket older than
 *                            that fails with GSPLAT_E_STATE.  A frame that overflowed the duplicate
 *                            buffer is re-run here (same camera) before the call returns.          */
#define GSPLAT_STREAM_SLOTS 3
int gsplat_render_streamed(gsplat_ctx *ctx, uint64_t *ticket);
int gsplat_acquire_frame(gsplat_ctx *ctx, uint64_t ticket, const uint8_t **bgr8, const uint32_t **counts);
/* Frames this context keeps in flight on the device (gsplat_config.concurrent_frames resolved).
 * A streaming caller gets full overlap with 2 tickets outstanding per frame in flight
 * (GSPLAT_STREAM_SLOTS - 1 each is the limit). */
int gsplat_concurrent_frames(const gsplat_ctx *ctx);

typedef struct gsplat_frame_stats {
  uint64_t num_gaussians; /* N                                             */
  uint64_t num_drawn;     /* V: Gaussians passing the render predicate      */
  uint64_t num_duplicates;/* D: (tile, Gaussian) pairs binned = sum of the list lengths  */
                          /* actually built (depth slices stop at saturated tiles)      */
  uint64_t dup_capacity;  /* current capacity of the duplicate buffers      */
  uint32_t max_tile_count;
  uint32_t num_tiles;     /* tiles_across * tiles_down                      */
  uint32_t regrown;       /* 1 if this frame had to be re-run after growing */
  uint32_t key_bits;      /* sorted key width: 32 depth bits + tile-id bits */
  uint64_t num_reference_duplicates; /* sum of the per-tile counts = D of the reference's lists */
  uint32_t depth_slices;  /* slices the frame was cut into                   */
  uint32_t num_depth_sorted; /* Gaussians that entered a depth sort (sort A), all slices */
} gsplat_frame_stats;
int gsplat_get_frame_stats(gsplat_ctx *ctx, gsplat_frame_stats *out);

/* Stage timings of the last frame in milliseconds (needs collect_timings):
 * [0] K1 project  [1] K3a depth sort  [2] K2 duplicate  [3] K3b tile sort  [4] K4 ranges
 * [5] K5 blend  [6] BGR8 post (last read_framebuffer_bgr8)  [7] whole frame.
 * Returns the count written. */
#define GSPLAT_NUM_TIMINGS 8
int gsplat_get_timings(gsplat_ctx *ctx, float *ms, int capacity);
/* Sums of the same GSPLAT_NUM_TIMINGS stage timings over all frames rendered since the last reset
 * (folded in without extra synchronisation while frames are being rendered), and the frame count. */
int gsplat_get_timings_accum(gsplat_ctx *ctx, double *ms_sum, uint64_t *frames, int reset);

/* ---- parity / debug taps (the reference exposes no equivalent; they exist so the stages can be
 * checked one by one against the oracle) ------------------------------------------------------ */
/* Projected records of the last frame (needs debug_records): rec8 = n x {px, py, clip_z, cov_a,
 * cov_b, cov_c, radius, drawn}; rect4 = n x {c0, c1, r0, r1} tile rectangle (valid iff drawn). */
int gsplat_read_projected(gsplat_ctx *ctx, float *rec8, int32_t *rect4);
/* Per-tile render lists of the last frame: offsets[num_tiles+1], lists[offsets[num_tiles]]
 * (Gaussian index = gid-1, in blend order).  lists may be NULL to query sizes only.  With more than
 * one depth slice the lists of all slices are only kept when debug_records is set; a tile's list is
 * then the reference's list cut after the slice in which the tile saturated. */
int gsplat_read_tile_lists(gsplat_ctx *ctx, int64_t *offsets, int32_t *lists, size_t lists_capacity);
/* Runs K2-K4 (and K5 when `blend` != 0) from caller-supplied projected inputs instead of K1's:
 * depth[n] = clip z, rect4 as above, drawn[n] (0/1).  K5 then uses the records of the last
 * gsplat_render().  "Per-tile lists bit-exact given identical projected inputs". */
int gsplat_bin_projected(gsplat_ctx *ctx, const float *depth, const int32_t *rect4,
                         const uint8_t *drawn, size_t n, int blend);
/* Stand-alone run of the K3 device radix sort on caller data: stable sort of (key, value) pairs
 * on key bits [begin_bit, end_bit) only, in ceil(bits/8) digits of equal width.  Used by the sort
 * known-answer tests. */
int gsplat_sort_pairs(gsplat_ctx *ctx, uint32_t *keys, uint32_t *values, size_t n, int begin_bit,
                      int end_bit);
/* The same sort without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Needs no gsplat_ctx.  The caller passes DEVICE arrays of
 * n words each (16-byte aligned, inputs distinct from outputs, on `device`); everything else the sort
 * needs lives in one file-scope device array of fixed size (8.5 MiB, held by every GPU that loads the
 * library until the process exits), so n is limited to GSPLAT_SORT_STATIC_MAX elements: a larger n
 * fails with GSPLAT_E_CAPACITY and a message in `err` (may be NULL).  Stable sort on key bits
 * [begin_bit, end_bit), asynchronous on `stream` (a cudaStream_t, NULL = default stream); calls no
 * cudaMalloc / cudaFree.  The array is shared by every call on a device and keeps the last call's
 * contents (the call clears what it needs cleared): calls that OVERLAP on the GPU race, serialise
 * them.  The first call on a device resolves the array's address (make it before a capture). */
#define GSPLAT_SORT_STATIC_MAX (1u << 20)
int gsplat_sort_pairs_static(int device, const uint32_t *d_keys_in, const uint32_t *d_values_in,
                             uint32_t *d_keys_out, uint32_t *d_values_out, size_t n, int begin_bit, int end_bit,
                             void *stream, char *err, size_t err_len);

/* ---- several GPUs of one process ---------------------------------------------
 * The reference drives ONE device (numIpus = 1, numReplicas = 1: src/main/splat.cpp:41-46); its unit
 * of parallelism is the framebuffer tile (one GSplat vertex per tile, ipu_rasteriser.cpp:334-385).
 * A group is the same IpuSplatter-shaped object over several B200s of this process:
 *   GSPLAT_GROUP_VIEWS  whole frames go to the member devices in turn (multi-view batches, orbits);
 *                       no inter-GPU traffic, each member streams its own frames to the host.
 *   GSPLAT_GROUP_BANDS  every frame is cut into tile-row bands, one per member (the scene is
 *                       replicated: it does not shard, transmittance is order dependent).  Each member
 *                       projects the whole scene and bins / sorts / blends its band; the BGR8 rows and
 *                       the band's tile counts are gathered on devices[0] with grouped ncclSend /
 *                       ncclRecv and delivered from there.  Cuts are re-balanced for every frame on the
 *                       per-tile counts (getIPUHistogram, ipu_rasteriser.cpp:104-106) of the newest
 *                       frame already acquired: no frame waits for the host.
 * cfg is the member configuration (cfg->device and the band fields are ignored).  The frame stream is
 * the context's: gsplat_group_render_streamed returns at once with a ticket, gsplat_group_acquire_frame
 * waits for that frame and hands out pointers into pinned host memory (valid until
 * gsplat_group_max_outstanding() further frames have been rendered).  NCCL is bound at run time
 * (libnccl.so.2); without it GSPLAT_GROUP_BANDS over more than one device fails with GSPLAT_E_NCCL. */
typedef struct gsplat_group gsplat_group;
#define GSPLAT_GROUP_VIEWS 0
#define GSPLAT_GROUP_BANDS 1
int gsplat_create_group(const gsplat_config *cfg, const int *devices, int n, int mode, gsplat_group **out);
void gsplat_destroy_group(gsplat_group *grp);
const char *gsplat_group_last_error(const gsplat_group *grp); /* grp may be NULL: last create error */
int gsplat_group_size(const gsplat_group *grp);
int gsplat_group_upload_scene(gsplat_group *grp, const float *gaussians16, size_t n);
int gsplat_group_set_camera(gsplat_group *grp, const float modelview[16], const float projection[16], float fov,
                            float scale_div);
int gsplat_group_render_streamed(gsplat_group *grp, uint64_t *ticket);
int gsplat_group_acquire_frame(gsplat_group *grp, uint64_t ticket, const uint8_t **bgr8, const uint32_t **counts);
int gsplat_group_max_outstanding(const gsplat_group *grp); /* tickets that may be rendered ahead of the acquired one */
int gsplat_group_sync(gsplat_group *grp);
/* GSPLAT_GROUP_BANDS: the n + 1 tile-row cuts the NEXT frame will use. Returns the count written. */
int gsplat_group_get_bands(const gsplat_group *grp, int *cuts, int capacity);
/* The member context of devices[index] (stats, parity taps; owned by the group). */
gsplat_ctx *gsplat_group_member(gsplat_group *grp, int index);
/* The band balancer by itself (host only): tile rows -> n_bands contiguous bands of nearly equal
 * predicted blend cost, cost of a tile = (count + 64) * tile_pixels; cuts[n_bands + 1]. */
int gsplat_balance_bands(const uint32_t *tile_counts, int tiles_across, int tiles_down, int tile_pixels, int n_bands,
                         int *cuts);

/* ---- helpers ------------------------------------------------------------------------------- */
/* Pinned host memory for frame read-back buffers (cudaHostAlloc / cudaFreeHost). */
int gsplat_host_alloc(void **ptr, size_t bytes);
int gsplat_host_free(void *ptr);
/* Device pointer of the BGR8 image / tiled f32 framebuffer of this context (for NCCL gathers of
 * bands / views by the caller's communicator; valid until destroy). */
int gsplat_device_framebuffer(gsplat_ctx *ctx, void **bgr8, void **tiled_f32);
/* Runs the BGR8 post-process on the device without copying to the host. */
int gsplat_postprocess_bgr8(gsplat_ctx *ctx);
/* Seeded synthetic scene of SURVEY §8d (xoshiro256**, Box-Muller), written as n Gaussian3D
 * records with the app's recentre + z-negate applied (src/main/splat.cpp:93-100). */
int gsplat_synth_scene(uint64_t seed, size_t n, float *gaussians16);
int gsplat_abi_version(void);
/* Device-time stopwatch on the context's stream (cudaEvent pair) for callers that time several
 * frames: begin records an event now, end records another, waits for it and returns the
 * milliseconds between them. */
int gsplat_timer_begin(gsplat_ctx *ctx);
int gsplat_timer_end(gsplat_ctx *ctx, float *ms);
/* Measurement hygiene for scenes smaller than the 126 MB L2: overwrites a 256 MB device scratch
 * buffer (allocated on first use) and waits, so the next frame starts with a cold L2. */
int gsplat_flush_l2(gsplat_ctx *ctx);
/* Kernel launches issued by this library since load (all contexts); for bench accounting. */
uint64_t gsplat_launch_count(void);
/* Re-runs K5 on the last frame's lists in counting mode: *pairs = number of (pixel, Gaussian)
 * pairs the blend loop visits (list entries per pixel up to and including the one that trips the
 * test_T < 1e-4 stop, codelets.cpp:405-408) — the unit the blend roofline is quoted in;
 * *blended = pairs that passed the alpha test and were accumulated. */
int gsplat_count_pairs(gsplat_ctx *ctx, uint64_t *pairs, uint64_t *blended);
/* Same counting run, plus per tile (num_tiles entries each, may be NULL): SM cycles the tile's CTA
 * took and the pairs it visited — the load-imbalance picture (reference README.md:16-17). */
int gsplat_profile_tiles(gsplat_ctx *ctx, uint64_t *pairs, uint64_t *blended, uint64_t *tile_cycles,
                         uint64_t *tile_pairs);

#ifdef __cplusplus
}
#endif
#endif /* GSPLAT_B200_H */
