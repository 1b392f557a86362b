"""Host-side mirror of the reference's ``splat::IpuSplatter`` (include/splat/ipu_rasteriser.hpp:20-55)
over the C ABI of ``libgsplat_b200.so`` (``include/gsplat_b200.h``).

The product is the shared library; this module is only the ctypes binding a Python caller (the
tests, ``bench.py``) uses — the same entry points a C++ host binds through
``include/gsplat_b200.hpp``.  There is no CPU path: if the library is missing or no CUDA device is
present, construction raises.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# GSPLAT_LIB: an experiment build of the same library (csrc/Makefile `variant`), for A/B measurements
LIB_PATH = os.environ.get("GSPLAT_LIB") or os.path.join(HERE, "libgsplat_b200.so")
CSRC = os.path.join(HERE, "csrc")

NUM_TIMINGS = 8
TIMING_NAMES = ("k1_project", "k3a_sort_depth", "k2_duplicate", "k3b_sort_tile", "k4_ranges", "k5_blend", "post_bgr8",
                "frame")


class GsplatError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"gsplat error {code}: {msg}")
        self.code = code


class Config(C.Structure):
    _fields_ = [("device", C.c_int), ("width", C.c_int), ("height", C.c_int), ("tile_w", C.c_int),
                ("tile_h", C.c_int), ("guard_factor", C.c_float), ("band_row_begin", C.c_int),
                ("band_row_end", C.c_int), ("collect_timings", C.c_int), ("debug_records", C.c_int),
                ("dup_capacity", C.c_size_t), ("depth_slices", C.c_int), ("concurrent_frames", C.c_int)]


class FrameStats(C.Structure):
    _fields_ = [("num_gaussians", C.c_uint64), ("num_drawn", C.c_uint64), ("num_duplicates", C.c_uint64),
                ("dup_capacity", C.c_uint64), ("max_tile_count", C.c_uint32), ("num_tiles", C.c_uint32),
                ("regrown", C.c_uint32), ("key_bits", C.c_uint32), ("num_reference_duplicates", C.c_uint64),
                ("depth_slices", C.c_uint32), ("num_depth_sorted", C.c_uint32)]


# every symbol include/gsplat_b200.h declares: name -> (restype, argtypes)
_vp, _fp = C.c_void_p, C.POINTER(C.c_float)
SYMBOLS = {
    "gsplat_abi_version": (C.c_int, []),
    "gsplat_config_default": (C.c_int, [C.POINTER(Config)]),
    "gsplat_create": (C.c_int, [C.POINTER(Config), C.POINTER(_vp)]),
    "gsplat_destroy": (None, [_vp]),
    "gsplat_last_error": (C.c_char_p, [_vp]),
    "gsplat_upload_scene": (C.c_int, [_vp, _fp, C.c_size_t]),
    "gsplat_set_camera": (C.c_int, [_vp, _fp, _fp, C.c_float, C.c_float]),
    "gsplat_set_band": (C.c_int, [_vp, C.c_int, C.c_int]),
    "gsplat_render": (C.c_int, [_vp]),
    "gsplat_sync": (C.c_int, [_vp]),
    "gsplat_render_streamed": (C.c_int, [_vp, C.POINTER(C.c_uint64)]),
    "gsplat_acquire_frame": (C.c_int, [_vp, C.c_uint64, C.POINTER(_vp), C.POINTER(_vp)]),
    "gsplat_read_framebuffer_tiled_f32": (C.c_int, [_vp, _vp]),
    "gsplat_read_framebuffer_bgr8": (C.c_int, [_vp, _vp, C.c_size_t]),
    "gsplat_read_tile_counts": (C.c_int, [_vp, _vp]),
    "gsplat_get_frame_stats": (C.c_int, [_vp, C.POINTER(FrameStats)]),
    "gsplat_get_timings": (C.c_int, [_vp, _fp, C.c_int]),
    "gsplat_read_projected": (C.c_int, [_vp, _vp, _vp]),
    "gsplat_read_tile_lists": (C.c_int, [_vp, _vp, _vp, C.c_size_t]),
    "gsplat_bin_projected": (C.c_int, [_vp, _vp, _vp, _vp, C.c_size_t, C.c_int]),
    "gsplat_sort_pairs": (C.c_int, [_vp, _vp, _vp, C.c_size_t, C.c_int, C.c_int]),
    "gsplat_sort_pairs_static": (C.c_int, [C.c_int, _vp, _vp, _vp, _vp, C.c_size_t, C.c_int, C.c_int, _vp,
                                           C.c_char_p, C.c_size_t]),
    "gsplat_host_alloc": (C.c_int, [C.POINTER(_vp), C.c_size_t]),
    "gsplat_host_free": (C.c_int, [_vp]),
    "gsplat_device_framebuffer": (C.c_int, [_vp, C.POINTER(_vp), C.POINTER(_vp)]),
    "gsplat_postprocess_bgr8": (C.c_int, [_vp]),
    "gsplat_synth_scene": (C.c_int, [C.c_uint64, C.c_size_t, _fp]),
    "gsplat_profile_tiles": (C.c_int, [_vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), _vp, _vp]),
    "gsplat_get_timings_accum": (C.c_int, [_vp, C.POINTER(C.c_double), C.POINTER(C.c_uint64), C.c_int]),
    "gsplat_timer_begin": (C.c_int, [_vp]),
    "gsplat_timer_end": (C.c_int, [_vp, _fp]),
    "gsplat_launch_count": (C.c_uint64, []),
    "gsplat_flush_l2": (C.c_int, [_vp]),
    "gsplat_create_group": (C.c_int, [C.POINTER(Config), C.POINTER(C.c_int), C.c_int, C.c_int, C.POINTER(_vp)]),
    "gsplat_destroy_group": (None, [_vp]),
    "gsplat_group_last_error": (C.c_char_p, [_vp]),
    "gsplat_group_size": (C.c_int, [_vp]),
    "gsplat_group_upload_scene": (C.c_int, [_vp, _fp, C.c_size_t]),
    "gsplat_group_set_camera": (C.c_int, [_vp, _fp, _fp, C.c_float, C.c_float]),
    "gsplat_group_render_streamed": (C.c_int, [_vp, C.POINTER(C.c_uint64)]),
    "gsplat_group_acquire_frame": (C.c_int, [_vp, C.c_uint64, C.POINTER(_vp), C.POINTER(_vp)]),
    "gsplat_group_max_outstanding": (C.c_int, [_vp]),
    "gsplat_group_sync": (C.c_int, [_vp]),
    "gsplat_group_get_bands": (C.c_int, [_vp, C.POINTER(C.c_int), C.c_int]),
    "gsplat_group_member": (_vp, [_vp, C.c_int]),
    "gsplat_balance_bands": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)]),
    "gsplat_concurrent_frames": (C.c_int, [_vp]),
    "gsplat_count_pairs": (C.c_int, [_vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
}

_lib = None


def build_library(force: bool = False) -> str:
    """Compile libgsplat_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "gsplat_b200.h")]
    stale = (not os.path.exists(LIB_PATH)) or any(os.path.getmtime(s) > os.path.getmtime(LIB_PATH)
                                                  for s in srcs if not s.endswith(".o"))
    if force or stale:
        if not os.path.exists("/usr/local/cuda/bin/nvcc") and os.path.exists(LIB_PATH):
            return LIB_PATH
        r = subprocess.run(["make", "-C", CSRC, "-j8"], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("building libgsplat_b200.so failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    return LIB_PATH


def load_library():
    """dlopen the library and bind every declared symbol.  Raises if it is missing (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(f"{LIB_PATH} is not built: run `make -C {CSRC}` (there is no CPU fallback)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def launch_count() -> int:
    return int(load_library().gsplat_launch_count())


def synth_scene(n: int, seed: int) -> np.ndarray:
    """Seeded synthetic scene as (n,16) Gaussian3D records; host-only call."""
    lib = load_library()
    out = np.empty((n, 16), np.float32)
    rc = lib.gsplat_synth_scene(seed, n, out.ctypes.data_as(_fp))
    if rc:
        raise GsplatError(rc, "gsplat_synth_scene")
    return out


class PinnedBuffer:
    """Page-locked host array (gsplat_host_alloc) for frame read-back."""

    def __init__(self, shape, dtype):
        self.lib = load_library()
        self.nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = _vp()
        rc = self.lib.gsplat_host_alloc(C.byref(p), self.nbytes)
        if rc:
            raise GsplatError(rc, "gsplat_host_alloc")
        self.ptr = p.value
        buf = (C.c_uint8 * self.nbytes).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=dtype).reshape(shape)

    def free(self):
        if self.ptr:
            self.array = None
            self.lib.gsplat_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class B200Splatter:
    """Same public surface as ``splat::IpuSplatter``: construct with the Gaussians and the tiled
    framebuffer geometry, ``updateModelView`` / ``updateProjection`` / ``updateFocalLengths`` per
    frame, ``execute`` (what ``GraphManager::execute`` triggers), ``getFrameBuffer``,
    ``getIPUHistogram``."""

    def __init__(self, gaussians16: np.ndarray, width=1280, height=720, tile_w=32, tile_h=20, device=0,
                 band=None, collect_timings=False, debug_records=False, dup_capacity=0, depth_slices=0,
                 concurrent_frames=0):
        self.lib = load_library()
        cfg = Config()
        self.lib.gsplat_config_default(C.byref(cfg))
        cfg.device, cfg.width, cfg.height, cfg.tile_w, cfg.tile_h = device, width, height, tile_w, tile_h
        if band is not None:
            cfg.band_row_begin, cfg.band_row_end = band
        cfg.collect_timings, cfg.debug_records, cfg.dup_capacity = int(collect_timings), int(debug_records), dup_capacity
        cfg.depth_slices = int(depth_slices)   # 0 = library default; 1 = complete (reference) lists
        cfg.concurrent_frames = int(concurrent_frames)   # 0 = library default (GSPLAT_DEFAULT_CONCURRENT_FRAMES = 4)
        self.cfg = cfg
        self.h = _vp()
        rc = self.lib.gsplat_create(C.byref(cfg), C.byref(self.h))
        if rc:
            self.h = _vp()
            raise GsplatError(rc, self.lib.gsplat_last_error(None).decode())
        self.width, self.height, self.tile_w, self.tile_h = width, height, tile_w, tile_h
        self.tiles_across, self.tiles_down = width // tile_w, height // tile_h
        self.num_tiles = self.tiles_across * self.tiles_down
        self._mv = np.eye(4, dtype=np.float32)
        self._proj = np.eye(4, dtype=np.float32)
        self._fxy = (float(np.radians(np.float32(40.0))), 0.1)
        self._slot_views = {}
        g = np.ascontiguousarray(gaussians16, dtype=np.float32).reshape(-1, 16)
        self.n = g.shape[0]
        self._check(self.lib.gsplat_upload_scene(self.h, g.ctypes.data_as(_fp), self.n))

    @property
    def concurrent_frames(self) -> int:
        """Frames the context keeps in flight on the device (lanes taking the frames round-robin)."""
        return int(self.lib.gsplat_concurrent_frames(self.h))

    # -- plumbing
    def _check(self, rc):
        if rc < 0:
            raise GsplatError(rc, self.lib.gsplat_last_error(self.h).decode())
        return rc

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            self.lib.gsplat_destroy(self.h)
            self.h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- IpuSplatter surface (ipu_rasteriser.cpp:86-110)
    def updateModelView(self, mv):
        self._mv = np.ascontiguousarray(mv, dtype=np.float32).reshape(4, 4)

    def updateProjection(self, mp):
        self._proj = np.ascontiguousarray(mp, dtype=np.float32).reshape(4, 4)

    def updateFocalLengths(self, fx, fy):
        self._fxy = (float(fx), float(fy))

    def set_band(self, row_begin, row_end):
        """Tile-row band [row_begin, row_end) rendered from the next execute() on (multi-GPU bands)."""
        self._check(self.lib.gsplat_set_band(self.h, int(row_begin), int(row_end)))

    def execute(self, sync=False):
        """One "project" program run (ipu_rasteriser.cpp:393-399,408-420); asynchronous."""
        self._check(self.lib.gsplat_set_camera(self.h, self._mv.ctypes.data_as(_fp), self._proj.ctypes.data_as(_fp),
                                               self._fxy[0], self._fxy[1]))
        self._check(self.lib.gsplat_render(self.h))
        if sync:
            self.sync()

    def sync(self):
        self._check(self.lib.gsplat_sync(self.h))

    def executeStreamed(self) -> int:
        """execute() whose BGR8 frame and tile counts are copied to a pinned host slot in the
        background (the render loop's imagePtr/imagePtrBuffered swap, src/main/splat.cpp:281-283).
        Returns the frame's ticket for acquireFrame()."""
        self._check(self.lib.gsplat_set_camera(self.h, self._mv.ctypes.data_as(_fp), self._proj.ctypes.data_as(_fp),
                                               self._fxy[0], self._fxy[1]))
        t = C.c_uint64()
        self._check(self.lib.gsplat_render_streamed(self.h, C.byref(t)))
        return int(t.value)

    def acquireFrame(self, ticket: int):
        """Waits for a streamed frame; returns ((H, W, 3) BGR8, (num_tiles,) counts) as views of the
        context's pinned slot, valid until the third-next executeStreamed()."""
        b, c = _vp(), _vp()
        self._check(self.lib.gsplat_acquire_frame(self.h, ticket, C.byref(b), C.byref(c)))
        views = self._slot_views.get(b.value)      # the slots are few and fixed: wrap each once
        if views is None:
            img = np.frombuffer((C.c_uint8 * (self.height * self.width * 3)).from_address(b.value), np.uint8)
            cnt = np.frombuffer((C.c_uint32 * self.num_tiles).from_address(c.value), np.uint32)
            views = self._slot_views[b.value] = (img.reshape(self.height, self.width, 3), cnt)
        return views

    def getFrameBuffer(self, out: np.ndarray | None = None) -> np.ndarray:
        """(H, W, 3) BGR8 image, ipu_rasteriser.cpp:131-144."""
        if out is None:
            out = np.empty((self.height, self.width, 3), np.uint8)
        self._check(self.lib.gsplat_read_framebuffer_bgr8(self.h, out.ctypes.data, out.strides[0]))
        return out

    def getIPUHistogram(self) -> np.ndarray:
        out = np.empty(self.num_tiles, np.uint32)
        self._check(self.lib.gsplat_read_tile_counts(self.h, out.ctypes.data))
        return out

    # -- beyond IpuSplatter: raw framebuffer, stats, parity taps
    def framebuffer_tiled(self, out: np.ndarray | None = None) -> np.ndarray:
        if out is None:
            out = np.empty(self.width * self.height * 4, np.float32)
        self._check(self.lib.gsplat_read_framebuffer_tiled_f32(self.h, out.ctypes.data))
        return out

    def stats(self) -> FrameStats:
        s = FrameStats()
        self._check(self.lib.gsplat_get_frame_stats(self.h, C.byref(s)))
        return s

    def timings(self) -> dict:
        ms = (C.c_float * NUM_TIMINGS)()
        self._check(self.lib.gsplat_get_timings(self.h, ms, NUM_TIMINGS))
        return dict(zip(TIMING_NAMES, [float(v) for v in ms]))

    def projected(self):
        rec = np.empty((self.n, 8), np.float32)
        rect = np.empty((self.n, 4), np.int32)
        self._check(self.lib.gsplat_read_projected(self.h, rec.ctypes.data, rect.ctypes.data))
        return rec, rect

    def tile_lists(self):
        offsets = np.empty(self.num_tiles + 1, np.int64)
        self._check(self.lib.gsplat_read_tile_lists(self.h, offsets.ctypes.data, None, 0))
        lists = np.empty(max(int(offsets[-1]), 1), np.int32)
        self._check(self.lib.gsplat_read_tile_lists(self.h, offsets.ctypes.data, lists.ctypes.data, lists.size))
        return offsets, lists[: int(offsets[-1])]

    def bin_projected(self, depth, rect4, drawn, blend=False):
        depth = np.ascontiguousarray(depth, np.float32)
        rect4 = np.ascontiguousarray(rect4, np.int32)
        drawn = np.ascontiguousarray(drawn, np.uint8)
        self._check(self.lib.gsplat_set_camera(self.h, self._mv.ctypes.data_as(_fp), self._proj.ctypes.data_as(_fp),
                                               self._fxy[0], self._fxy[1]))
        self._check(self.lib.gsplat_bin_projected(self.h, depth.ctypes.data, rect4.ctypes.data, drawn.ctypes.data,
                                                  depth.size, int(blend)))

    def sort_pairs(self, keys, values, begin_bit=0, end_bit=32):
        keys = np.ascontiguousarray(keys, np.uint32).copy()
        values = np.ascontiguousarray(values, np.uint32).copy()
        self._check(self.lib.gsplat_sort_pairs(self.h, keys.ctypes.data, values.ctypes.data, keys.size, begin_bit,
                                               end_bit))
        return keys, values

    def device_pointers(self):
        b, f = _vp(), _vp()
        self._check(self.lib.gsplat_device_framebuffer(self.h, C.byref(b), C.byref(f)))
        return b.value, f.value

    def timings_accum(self, reset=False):
        """(mean ms per stage over the frames since the last reset, frame count)."""
        ms = (C.c_double * NUM_TIMINGS)()
        n = C.c_uint64()
        self._check(self.lib.gsplat_get_timings_accum(self.h, ms, C.byref(n), int(reset)))
        k = max(int(n.value), 1)
        return dict(zip(TIMING_NAMES, [float(v) / k for v in ms])), int(n.value)

    def timer_begin(self):
        self._check(self.lib.gsplat_timer_begin(self.h))

    def timer_end(self) -> float:
        ms = C.c_float()
        self._check(self.lib.gsplat_timer_end(self.h, C.byref(ms)))
        return float(ms.value)

    def count_pairs(self):
        a, b = C.c_uint64(), C.c_uint64()
        self._check(self.lib.gsplat_count_pairs(self.h, C.byref(a), C.byref(b)))
        return int(a.value), int(b.value)

    def profile_tiles(self):
        """(pairs, blended, per-tile SM cycles, per-tile visited pairs) of a counting K5 re-run."""
        a, b = C.c_uint64(), C.c_uint64()
        cyc = np.zeros(self.num_tiles, np.uint64)
        prs = np.zeros(self.num_tiles, np.uint64)
        self._check(self.lib.gsplat_profile_tiles(self.h, C.byref(a), C.byref(b), cyc.ctypes.data, prs.ctypes.data))
        return int(a.value), int(b.value), cyc, prs

    def flush_l2(self):
        """Cold L2 before the next frame (scenes smaller than L2): 256 MB device write, synchronous."""
        self._check(self.lib.gsplat_flush_l2(self.h))

    def postprocess_bgr8(self):
        self._check(self.lib.gsplat_postprocess_bgr8(self.h))


SORT_STATIC_MAX = 1 << 20   # GSPLAT_SORT_STATIC_MAX (include/gsplat_b200.h)


def sort_pairs_static(device: int, d_keys_in: int, d_values_in: int, d_keys_out: int, d_values_out: int, n: int,
                      begin_bit: int = 0, end_bit: int = 32, stream: int = 0) -> None:
    """gsplat_sort_pairs_static: the device radix sort on the caller's DEVICE arrays (addresses as
    ints) with its scratch in a fixed file-scope device array; allocates nothing, asynchronous on
    `stream`.  Raises GsplatError with the library's message (e.g. n > SORT_STATIC_MAX)."""
    err = C.create_string_buffer(256)
    rc = load_library().gsplat_sort_pairs_static(device, d_keys_in, d_values_in, d_keys_out, d_values_out, n,
                                                 begin_bit, end_bit, stream, err, len(err))
    if rc != 0:
        raise GsplatError(rc, err.value.decode())


GROUP_VIEWS, GROUP_BANDS = 0, 1


def balance_bands_native(tile_counts: np.ndarray, tiles_across: int, tiles_down: int, n_bands: int, tile_pixels: int = 640):
    """The library's band balancer (gsplat_balance_bands; host only): n_bands + 1 tile-row cuts."""
    lib = load_library()
    c = np.ascontiguousarray(tile_counts, np.uint32)
    cuts = (C.c_int * (n_bands + 1))()
    rc = lib.gsplat_balance_bands(c.ctypes.data, tiles_across, tiles_down, tile_pixels, n_bands, cuts)
    if rc:
        raise GsplatError(rc, "gsplat_balance_bands")
    return list(cuts)


class B200Group:
    """Several B200s of this process behind one IpuSplatter-shaped object (gsplat_create_group):
    mode GROUP_VIEWS (whole frames to the devices in turn) or GROUP_BANDS (every frame cut into
    tile-row bands, gathered on devices[0] over NCCL).  The reference drives one IPU
    (src/main/splat.cpp:41-46); the surface mirrored here is the same updateModelView /
    updateProjection / updateFocalLengths + the frame stream."""

    def __init__(self, gaussians16, devices, mode=GROUP_BANDS, width=1280, height=720, tile_w=32, tile_h=20,
                 depth_slices=0, concurrent_frames=0, dup_capacity=0):
        self.lib = load_library()
        cfg = Config()
        self.lib.gsplat_config_default(C.byref(cfg))
        cfg.width, cfg.height, cfg.tile_w, cfg.tile_h = width, height, tile_w, tile_h
        cfg.depth_slices, cfg.concurrent_frames, cfg.dup_capacity = int(depth_slices), int(concurrent_frames), dup_capacity
        devs = (C.c_int * len(devices))(*devices)
        self.h = _vp()
        rc = self.lib.gsplat_create_group(C.byref(cfg), devs, len(devices), int(mode), C.byref(self.h))
        if rc:
            self.h = _vp()
            raise GsplatError(rc, self.lib.gsplat_group_last_error(None).decode())
        self.mode, self.n = mode, len(devices)
        self.width, self.height, self.tile_w, self.tile_h = width, height, tile_w, tile_h
        self.tiles_across, self.tiles_down = width // tile_w, height // tile_h
        self.num_tiles = self.tiles_across * self.tiles_down
        self._mv = np.eye(4, dtype=np.float32)
        self._proj = np.eye(4, dtype=np.float32)
        self._fxy = (float(np.radians(np.float32(40.0))), 0.1)
        self._slot_views = {}
        g = np.ascontiguousarray(gaussians16, dtype=np.float32).reshape(-1, 16)
        self._check(self.lib.gsplat_group_upload_scene(self.h, g.ctypes.data_as(_fp), g.shape[0]))

    def _check(self, rc):
        if rc < 0:
            raise GsplatError(rc, self.lib.gsplat_group_last_error(self.h).decode())
        return rc

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            self.lib.gsplat_destroy_group(self.h)
            self.h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def updateModelView(self, mv):
        self._mv = np.ascontiguousarray(mv, dtype=np.float32).reshape(4, 4)

    def updateProjection(self, mp):
        self._proj = np.ascontiguousarray(mp, dtype=np.float32).reshape(4, 4)

    def updateFocalLengths(self, fx, fy):
        self._fxy = (float(fx), float(fy))

    @property
    def max_outstanding(self) -> int:
        return int(self.lib.gsplat_group_max_outstanding(self.h))

    def bands(self):
        cuts = (C.c_int * (self.n + 1))()
        self._check(self.lib.gsplat_group_get_bands(self.h, cuts, self.n + 1))
        return list(cuts)

    def executeStreamed(self) -> int:
        self._check(self.lib.gsplat_group_set_camera(self.h, self._mv.ctypes.data_as(_fp), self._proj.ctypes.data_as(_fp),
                                                     self._fxy[0], self._fxy[1]))
        t = C.c_uint64()
        self._check(self.lib.gsplat_group_render_streamed(self.h, C.byref(t)))
        return int(t.value)

    def acquireFrame(self, ticket: int):
        b, c = _vp(), _vp()
        self._check(self.lib.gsplat_group_acquire_frame(self.h, ticket, C.byref(b), C.byref(c)))
        views = self._slot_views.get(b.value)
        if views is None:
            img = np.frombuffer((C.c_uint8 * (self.height * self.width * 3)).from_address(b.value), np.uint8)
            cnt = np.frombuffer((C.c_uint32 * self.num_tiles).from_address(c.value), np.uint32)
            views = self._slot_views[b.value] = (img.reshape(self.height, self.width, 3), cnt)
        return views

    def sync(self):
        self._check(self.lib.gsplat_group_sync(self.h))
