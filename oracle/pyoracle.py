"""ctypes bindings for the CPU oracle — TEST INFRASTRUCTURE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import this module.  Two back ends with the same interface:

* ``kind="reference"`` — ``oracle/_ref/libsplat_ref.so``: the reference's own code
  (codelets/splat/codelets.cpp + include/splat/*.hpp), compiled by ``oracle/build_ref.sh``.
* ``kind="port"``      — ``oracle/liboracle.so``: the plain-C restatement ``splat_oracle.c``.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libsplat_ref.so")
PORT_SO = os.path.join(HERE, "liboracle.so")
REF_DATA = os.path.join(HERE, "_ref", "data")

_fp = C.POINTER(C.c_float)


def build_port() -> str:
    if not os.path.exists(PORT_SO) or os.path.getmtime(PORT_SO) < os.path.getmtime(os.path.join(HERE, "splat_oracle.c")):
        subprocess.run(["make", "-C", HERE, "liboracle.so"], check=True, capture_output=True)
    return PORT_SO


def build_ref() -> str | None:
    """Build oracle/_ref from /root/reference when it is present (this container only)."""
    if os.path.isdir("/root/reference/codelets"):
        subprocess.run([os.path.join(HERE, "build_ref.sh")], check=True, capture_output=True)
    return REF_SO if os.path.exists(REF_SO) else None


def have_ref() -> bool:
    return os.path.exists(REF_SO)


@dataclass
class Frame:
    rec: np.ndarray        # (n,8) px,py,clip_z,cov_a,cov_b,cov_c,radius,drawn
    rect: np.ndarray       # (n,4) c0,c1,r0,r1
    counts: np.ndarray     # (tiles,) u32
    offsets: np.ndarray    # (tiles+1,) i64
    lists: np.ndarray      # (D,) i32 gaussian index, per tile in (z asc, index asc) order
    fb: np.ndarray | None  # (W*H*4,) f32 tile-major
    ms: tuple              # project, bin, blend (CPU milliseconds)
    pairs: int = 0

    @property
    def drawn(self) -> np.ndarray:
        return self.rec[:, 7] > 0


class Oracle:
    def __init__(self, kind: str = "port", tie_order: str = "canonical"):
        """tie_order: how Gaussians with bit-identical clip z are ordered in the blend — "canonical"
        (z, then gid: the parity contract) or "quicksort" (gid-order fill, the
        reference's unstable quicksort decides; the lists are reported canonically either way)."""
        if kind == "reference":
            if not os.path.exists(REF_SO):
                raise FileNotFoundError(REF_SO)
            path, self.p = REF_SO, "ref_"
        elif kind == "port":
            path, self.p = build_port(), "orc_"
        else:
            raise ValueError(kind)
        self.kind = kind
        self.lib = C.CDLL(path)
        f = self._fn
        f("frame_create", C.c_void_p, [_fp, C.c_size_t, _fp, _fp, C.c_float, C.c_float] + [C.c_int] * 6)
        f("frame_destroy", None, [C.c_void_p])
        f("frame_num_duplicates", C.c_size_t, [C.c_void_p])
        f("frame_get_records", None, [C.c_void_p, _fp, C.POINTER(C.c_int32)])
        f("frame_get_lists", None, [C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_int64), C.POINTER(C.c_int32)])
        f("frame_get_framebuffer", None, [C.c_void_p, _fp])
        f("frame_get_timings", None, [C.c_void_p, C.POINTER(C.c_double)])
        f("num_threads", C.c_int, [])
        f("set_tie_order", None, [C.c_int])
        if tie_order not in ("canonical", "quicksort"):
            raise ValueError(tie_order)
        self.tie_order = tie_order
        f("conic_opacity", None, [_fp, C.c_float, _fp])
        f("tile_bounds", None, [C.c_int] * 4 + [C.c_uint, _fp])
        if kind == "reference":   # O1: the unmodified GSplat::compute over the 4-neighbour lattice
            f("lattice_create", C.c_void_p, [_fp, C.c_size_t] + [C.c_int] * 4 + [C.c_size_t] * 2)
            f("lattice_step", None, [C.c_void_p, _fp, _fp, C.c_float, C.c_float, C.c_int])
            f("lattice_get", None, [C.c_void_p, C.POINTER(C.c_uint32), _fp, C.POINTER(C.c_double)])
            f("lattice_rendered", C.c_size_t, [C.c_void_p, C.c_int, _fp, C.c_size_t])
            f("lattice_destroy", None, [C.c_void_p])
        if kind == "port":
            f("frame_num_pairs", C.c_uint64, [C.c_void_p])
            f("framebuffer_to_bgr8", None, [_fp] + [C.c_int] * 4 + [C.POINTER(C.c_uint8)])

    def _fn(self, name, res, args):
        fn = getattr(self.lib, self.p + name)
        fn.restype, fn.argtypes = res, args
        setattr(self, "_" + name, fn)

    @property
    def cores(self) -> int:
        return int(self._num_threads())

    def frame(self, g16, view, proj, fov, scale_div, W, H, tile_w=32, tile_h=20, render=True, threads=0) -> Frame:
        g16 = np.ascontiguousarray(g16, dtype=np.float32).reshape(-1, 16)
        view = np.ascontiguousarray(view, dtype=np.float32).ravel()
        proj = np.ascontiguousarray(proj, dtype=np.float32).ravel()
        n = g16.shape[0]
        TA, TD = W // tile_w, H // tile_h
        assert TA * tile_w == W and TD * tile_h == H
        self._set_tie_order(int(self.tie_order == "quicksort"))   # library-global: set per call
        h = self._frame_create(g16.ctypes.data_as(_fp), n, view.ctypes.data_as(_fp), proj.ctypes.data_as(_fp),
                               fov, scale_div, W, H, TA, TD, int(render), threads)
        try:
            D = int(self._frame_num_duplicates(h))
            rec = np.empty((n, 8), np.float32)
            rect = np.empty((n, 4), np.int32)
            self._frame_get_records(h, rec.ctypes.data_as(_fp), rect.ctypes.data_as(C.POINTER(C.c_int32)))
            counts = np.empty(TA * TD, np.uint32)
            offsets = np.empty(TA * TD + 1, np.int64)
            lists = np.empty(max(D, 1), np.int32)
            self._frame_get_lists(h, counts.ctypes.data_as(C.POINTER(C.c_uint32)),
                                  offsets.ctypes.data_as(C.POINTER(C.c_int64)),
                                  lists.ctypes.data_as(C.POINTER(C.c_int32)))
            fb = None
            if render:
                fb = np.empty(W * H * 4, np.float32)
                self._frame_get_framebuffer(h, fb.ctypes.data_as(_fp))
            ms = (C.c_double * 3)()
            self._frame_get_timings(h, ms)
            pairs = int(self._frame_num_pairs(h)) if self.kind == "port" else 0
        finally:
            self._frame_destroy(h)
        return Frame(rec, rect, counts, offsets, lists[:D], fb, tuple(ms), pairs)

    def lattice(self, g16, view, proj, fov, scale_div, W, H, frames, spare_slots=600, chan_slots=75,
                tile_w=32, tile_h=20, threads=0):
        """O1 (kind="reference" only): run the reference's own GSplat vertices + edge lattice for
        `frames` program runs (ipu_rasteriser.cpp:393-399); defaults = the as-shipped capacities
        (ipu_rasteriser.cpp:307-316).  Returns (counts, fb_tiled, per-frame total counts,
        per-tile rendered lists as (k,3) arrays of (mean.x, mean.y, z) in blend order)."""
        assert self.kind == "reference"
        g16 = np.ascontiguousarray(g16, dtype=np.float32).reshape(-1, 16)
        view = np.ascontiguousarray(view, dtype=np.float32).ravel()
        proj = np.ascontiguousarray(proj, dtype=np.float32).ravel()
        TA, TD = W // tile_w, H // tile_h
        h = self._lattice_create(g16.ctypes.data_as(_fp), g16.shape[0], W, H, TA, TD, spare_slots, chan_slots)
        try:
            counts = np.zeros(TA * TD, np.uint32)
            fb = np.zeros(W * H * 4, np.float32)
            ms = (C.c_double * 2)()
            totals = []
            for _ in range(frames):
                self._lattice_step(h, view.ctypes.data_as(_fp), proj.ctypes.data_as(_fp), fov, scale_div, threads)
                self._lattice_get(h, counts.ctypes.data_as(C.POINTER(C.c_uint32)), fb.ctypes.data_as(_fp), ms)
                totals.append(int(counts.sum()))
            held = []
            buf = np.empty((g16.shape[0] + 1, 3), np.float32)
            for t in range(TA * TD):
                k = int(self._lattice_rendered(h, t, buf.ctypes.data_as(_fp), buf.shape[0])) if counts[t] else 0
                held.append(buf[:k].copy())
        finally:
            self._lattice_destroy(h)
        return counts, fb, totals, held

    def lattice_time(self, g16, view, proj, fov, scale_div, W, H, warm=60, timed=5, spare_slots=600, chan_slots=75,
                     tile_w=32, tile_h=20, threads=0):
        """O1 "as shipped" timing (kind="reference" only): `warm` program runs so the lattice reaches
        its (lossy) steady state, then `timed` runs on the clock.  Returns (seconds per frame,
        rendered copies in the last frame)."""
        import time
        assert self.kind == "reference"
        g16 = np.ascontiguousarray(g16, dtype=np.float32).reshape(-1, 16)
        view = np.ascontiguousarray(view, dtype=np.float32).ravel()
        proj = np.ascontiguousarray(proj, dtype=np.float32).ravel()
        TA, TD = W // tile_w, H // tile_h
        h = self._lattice_create(g16.ctypes.data_as(_fp), g16.shape[0], W, H, TA, TD, spare_slots, chan_slots)
        try:
            step = lambda: self._lattice_step(h, view.ctypes.data_as(_fp), proj.ctypes.data_as(_fp), fov, scale_div, threads)
            for _ in range(warm):
                step()
            t0 = time.perf_counter()
            for _ in range(timed):
                step()
            dt = (time.perf_counter() - t0) / max(timed, 1)
            counts = np.zeros(TA * TD, np.uint32)
            fb = np.zeros(W * H * 4, np.float32)
            ms = (C.c_double * 2)()
            self._lattice_get(h, counts.ctypes.data_as(C.POINTER(C.c_uint32)), fb.ctypes.data_as(_fp), ms)
        finally:
            self._lattice_destroy(h)
        return dt, int(counts.sum())

    def conic_opacity(self, cov, opacity):
        cov = np.ascontiguousarray(cov, np.float32)
        out = np.empty(4, np.float32)
        self._conic_opacity(cov.ctypes.data_as(_fp), opacity, out.ctypes.data_as(_fp))
        return out

    def tile_bounds(self, W, H, TA, TD, tid):
        out = np.empty(4, np.float32)
        self._tile_bounds(W, H, TA, TD, tid, out.ctypes.data_as(_fp))
        return out

    def to_bgr8(self, fb, W, H, tile_w=32, tile_h=20):
        assert self.kind == "port"
        fb = np.ascontiguousarray(fb, np.float32)
        out = np.empty((H, W, 3), np.uint8)
        self._framebuffer_to_bgr8(fb.ctypes.data_as(_fp), W, H, W // tile_w, H // tile_h,
                                  out.ctypes.data_as(C.POINTER(C.c_uint8)))
        return out


def detile(fb_tiled: np.ndarray, W: int, H: int, tile_w=32, tile_h=20) -> np.ndarray:
    """tile-major f32 chunks -> (H, W, 4) image (layout of ipu_rasteriser.cpp:115-129)."""
    TA, TD = W // tile_w, H // tile_h
    return (fb_tiled.reshape(TD, TA, tile_h, tile_w, 4).transpose(0, 2, 1, 3, 4).reshape(H, W, 4))


def psnr(a: np.ndarray, b: np.ndarray, peak: float = 1.0) -> float:
    mse = float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2))
    return float("inf") if mse == 0 else 10.0 * np.log10(peak * peak / mse)


# ---- the reference's host-side INPUT code (oracle/_ref/libhost_ref.so) --------------------------------
HOST_SO = os.path.join(HERE, "_ref", "libhost_ref.so")


def have_host_ref() -> bool:
    return os.path.exists(HOST_SO)


class HostRef:
    """The reference's own PLY loading (happly via src/splat/file_io.cpp), lookAtBoundingBox
    (src/splat/camera.cpp), fitFrustumToBoundingBox (src/splat/geometry.cpp) and glm, compiled
    unchanged (oracle/build_ref.sh) behind oracle/ref_host_driver.cpp.  Pins scene.py and the C++
    host driver to the reference.  TEST INFRASTRUCTURE."""

    def __init__(self):
        if not os.path.exists(HOST_SO):
            raise FileNotFoundError(HOST_SO)
        lib = self.lib = C.CDLL(HOST_SO)
        lib.refhost_open.restype = C.c_void_p
        lib.refhost_open.argtypes = [C.c_char_p]
        lib.refhost_error.restype = C.c_char_p
        lib.refhost_error.argtypes = [C.c_void_p]
        lib.refhost_close.argtypes = [C.c_void_p]
        lib.refhost_count.restype = C.c_size_t
        lib.refhost_count.argtypes = [C.c_void_p]
        lib.refhost_properties.argtypes = [C.c_void_p, _fp]
        lib.refhost_gaussians.argtypes = [C.c_void_p, _fp]
        lib.refhost_bounds.argtypes = [C.c_void_p, _fp, _fp]
        lib.refhost_look_at_bounding_box.argtypes = [_fp, _fp, _fp, C.c_float, _fp]
        lib.refhost_fit_frustum.argtypes = [_fp, _fp, C.c_float, C.c_float, _fp]
        lib.refhost_app_projection.argtypes = [_fp, _fp, C.c_float, C.c_float, _fp]
        lib.refhost_orbit_view.argtypes = [C.c_float, _fp]
        lib.refhost_ui_view.argtypes = [_fp] + [C.c_float] * 5 + [_fp]
        lib.refhost_rotate.argtypes = [_fp, C.c_float, _fp, _fp]

    @staticmethod
    def _p(a):
        return np.ascontiguousarray(a, np.float32).ctypes.data_as(_fp)

    def load(self, path: str):
        """(properties n x 14 in file_io.cpp:62-77 order, Gaussian3D records n x 16, bbox min, bbox max)
        after the app's recentre + z-negate."""
        h = self.lib.refhost_open(path.encode())
        try:
            err = self.lib.refhost_error(h)
            if err:
                raise RuntimeError(err.decode())
            n = int(self.lib.refhost_count(h))
            props = np.empty((n, 14), np.float32)
            g16 = np.empty((n, 16), np.float32)
            bmin, bmax = np.empty(3, np.float32), np.empty(3, np.float32)
            self.lib.refhost_properties(h, props.ctypes.data_as(_fp))
            self.lib.refhost_gaussians(h, g16.ctypes.data_as(_fp))
            self.lib.refhost_bounds(h, bmin.ctypes.data_as(_fp), bmax.ctypes.data_as(_fp))
        finally:
            self.lib.refhost_close(h)
        return props, g16, bmin, bmax

    def _mat(self, fn, *args):
        out = np.empty(16, np.float32)
        fn(*args, out.ctypes.data_as(_fp))
        return out.reshape(4, 4)

    def look_at_bounding_box(self, bmin, bmax, up, scale):
        a, b, u = (np.ascontiguousarray(x, np.float32) for x in (bmin, bmax, up))
        return self._mat(self.lib.refhost_look_at_bounding_box, a.ctypes.data_as(_fp), b.ctypes.data_as(_fp),
                         u.ctypes.data_as(_fp), float(scale))

    def fit_frustum(self, bmin, bmax, fov, aspect):
        a, b = (np.ascontiguousarray(x, np.float32) for x in (bmin, bmax))
        return self._mat(self.lib.refhost_fit_frustum, a.ctypes.data_as(_fp), b.ctypes.data_as(_fp), float(fov), float(aspect))

    def app_projection(self, bmin, bmax, fov, aspect):
        a, b = (np.ascontiguousarray(x, np.float32) for x in (bmin, bmax))
        return self._mat(self.lib.refhost_app_projection, a.ctypes.data_as(_fp), b.ctypes.data_as(_fp), float(fov), float(aspect))

    def orbit_view(self, degrees):
        return self._mat(self.lib.refhost_orbit_view, float(degrees))

    def ui_view(self, model_view, env_deg, env_deg2, X, Y, Z):
        m = np.ascontiguousarray(model_view, np.float32)
        return self._mat(self.lib.refhost_ui_view, m.ctypes.data_as(_fp), float(env_deg), float(env_deg2), float(X), float(Y), float(Z))

    def rotate(self, m, angle, axis):
        mm, ax = np.ascontiguousarray(m, np.float32), np.ascontiguousarray(axis, np.float32)
        return self._mat(self.lib.refhost_rotate, mm.ctypes.data_as(_fp), float(angle), ax.ctypes.data_as(_fp))
