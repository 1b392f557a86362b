"""Scene and camera inputs of the splat hot path (host side, numpy f32).

These are the *inputs* the reference's render loop feeds to ``IpuSplatter`` — they are not part
of the accelerated path, but tests and ``bench.py`` need them to drive it the way
``src/main/splat.cpp`` does:

* ``load_ply``                — the 14 float properties ``src/splat/file_io.cpp:62-77`` pulls
* ``gaussians_from_ply``      — PLY -> ``Gaussian3D`` records, ``src/main/splat.cpp:93-100,136-163``
* ``look_at_bounding_box``    — ``src/splat/camera.cpp:10-15``
* ``fit_frustum_to_bounding_box`` — ``src/splat/geometry.cpp:9-24``
* ``mvp_start`` / ``orbit_view``  — ``src/main/splat.cpp:235-241`` and SURVEY.md §8d config 2
* ``ui_view``                 — ``src/main/splat.cpp:312-314``

Matrices are ``(4, 4)`` float32 arrays in glm memory order: ``m[col, row]``; ``m.ravel()`` is what
``glm::value_ptr`` yields and what the C ABI takes.
"""
from __future__ import annotations

import numpy as np

SH_C0 = np.float32(0.28209479177387814)  # src/main/splat.cpp:136
F32 = np.float32

PLY_PROPS = ("x", "y", "z", "f_dc_0", "f_dc_1", "f_dc_2", "opacity",
             "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3")


def load_ply(path: str) -> dict[str, np.ndarray]:
    """Binary-little-endian PLY with float vertex properties (what happly reads for the
    reference, ``src/splat/file_io.cpp:30-42,62-77``)."""
    with open(path, "rb") as fh:
        names, n, fmt = [], 0, None
        while True:
            line = fh.readline()
            if not line:
                raise ValueError(f"{path}: no end_header")
            tok = line.decode("ascii", "replace").split()
            if not tok:
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                if tok[1] != "vertex" and names:
                    break
                if tok[1] == "vertex":
                    n = int(tok[2])
            elif tok[0] == "property":
                if tok[1] != "float":
                    raise ValueError(f"{path}: non-float property {tok}")
                names.append(tok[2])
            elif tok[0] == "end_header":
                break
        if fmt != "binary_little_endian":
            raise ValueError(f"{path}: unsupported PLY format {fmt}")
        raw = np.frombuffer(fh.read(n * len(names) * 4), dtype="<f4").reshape(n, len(names))
    missing = [p for p in PLY_PROPS if p not in names]
    if missing:
        raise ValueError(f"{path}: missing properties {missing}")
    return {p: np.ascontiguousarray(raw[:, names.index(p)]) for p in PLY_PROPS}


def pack_gaussians(pos, rgb, opacity, rot, log_scale) -> np.ndarray:
    """(N,16) f32 ``Gaussian3D`` records: mean4 colour4 rot4 scale3 gid
    (``include/splat/ipu_geometry.hpp:305-311``); gid = i+1 (``src/main/splat.cpp:161``)."""
    n = pos.shape[0]
    g = np.zeros((n, 16), dtype=F32)
    g[:, 0:3] = pos
    g[:, 3] = 1.0
    g[:, 4:7] = rgb
    g[:, 7] = opacity
    g[:, 8:12] = rot
    g[:, 12:15] = log_scale
    g[:, 15] = np.arange(1, n + 1, dtype=F32)
    return g


def recentre_negate_z(pos: np.ndarray) -> np.ndarray:
    """``src/main/splat.cpp:93-100``: translate so the bbox centroid is zero, then negate z."""
    pos = pos.astype(F32, copy=True)
    centre = (pos.max(axis=0) + pos.min(axis=0)) * F32(0.5)
    pos -= centre
    pos[:, 2] = -pos[:, 2]
    return pos


def gaussians_from_ply(ply: dict[str, np.ndarray]) -> np.ndarray:
    pos = recentre_negate_z(np.stack([ply["x"], ply["y"], ply["z"]], axis=1))
    f_dc = np.stack([ply["f_dc_0"], ply["f_dc_1"], ply["f_dc_2"]], axis=1).astype(F32)
    rgb = np.maximum(SH_C0 * f_dc + F32(0.5), F32(0.0))            # splat.cpp:143-148 (no upper clamp)
    rot = np.stack([ply[f"rot_{k}"] for k in range(4)], axis=1)     # rot_0 = quaternion w
    scale = np.stack([ply[f"scale_{k}"] for k in range(3)], axis=1)  # raw log-scales
    return pack_gaussians(pos, rgb, ply["opacity"], rot, scale)    # raw (logit) opacity


# ----------------------------------------------------------------------------- glm restatements
# glm::tan / cos / sin on floats are the C library's tanf / cosf / sinf: call those (numpy's SIMD
# kernels round the last bit differently for some arguments), so the matrices below are bit-identical
# to the ones the reference's host code builds (tests/test_host_inputs.py).
def _libm_f32(name, fallback):
    try:
        import ctypes
        fn = getattr(ctypes.CDLL("libm.so.6"), name)
        fn.restype, fn.argtypes = ctypes.c_float, [ctypes.c_float]
        return lambda x: F32(fn(float(F32(x))))
    except (OSError, AttributeError):
        return lambda x: fallback(F32(x), dtype=F32)


_tanf, _cosf, _sinf = _libm_f32("tanf", np.tan), _libm_f32("cosf", np.cos), _libm_f32("sinf", np.sin)


def _dot3(a, b):
    """glm::dot for vec3 (detail/func_geometric.inl): tmp = a * b; tmp.x + tmp.y + tmp.z, every step
    rounded to f32 (np.dot goes through BLAS, which may fuse or reorder)."""
    a, b = np.asarray(a, dtype=F32), np.asarray(b, dtype=F32)
    t = a * b
    return F32(F32(t[0] + t[1]) + t[2])


def _normalize(v):
    v = np.asarray(v, dtype=F32)
    return v * (F32(1.0) / np.sqrt(_dot3(v, v), dtype=F32))


def look_at(eye, center, up) -> np.ndarray:
    """glm::lookAtRH (ext/matrix_transform.inl:153-173)."""
    eye, center, up = (np.asarray(a, dtype=F32) for a in (eye, center, up))
    f = _normalize(center - eye)
    s = _normalize(np.cross(f, up).astype(F32))
    u = np.cross(s, f).astype(F32)
    m = np.eye(4, dtype=F32)
    m[0, 0], m[1, 0], m[2, 0] = s
    m[0, 1], m[1, 1], m[2, 1] = u
    m[0, 2], m[1, 2], m[2, 2] = -f
    m[3, 0] = -_dot3(s, eye)
    m[3, 1] = -_dot3(u, eye)
    m[3, 2] = _dot3(f, eye)
    return m


def frustum(left, right, bottom, top, near, far) -> np.ndarray:
    """glm::frustumRH_NO (ext/matrix_clip_space.inl:165-176)."""
    left, right, bottom, top, near, far = (F32(a) for a in (left, right, bottom, top, near, far))
    m = np.zeros((4, 4), dtype=F32)
    m[0, 0] = (F32(2) * near) / (right - left)
    m[1, 1] = (F32(2) * near) / (top - bottom)
    m[2, 0] = (right + left) / (right - left)
    m[2, 1] = (top + bottom) / (top - bottom)
    m[2, 2] = -(far + near) / (far - near)
    m[2, 3] = F32(-1)
    m[3, 2] = -(F32(2) * far * near) / (far - near)
    return m


def rotate(m: np.ndarray, angle: float, axis) -> np.ndarray:
    """glm::rotate(m, angle, axis) (ext/matrix_transform.inl:18-46)."""
    a = F32(angle)
    c, s = _cosf(a), _sinf(a)
    ax = _normalize(axis)
    t = (F32(1) - c) * ax
    r = np.zeros((3, 3), dtype=F32)
    r[0, 0] = c + t[0] * ax[0]; r[0, 1] = t[0] * ax[1] + s * ax[2]; r[0, 2] = t[0] * ax[2] - s * ax[1]
    r[1, 0] = t[1] * ax[0] - s * ax[2]; r[1, 1] = c + t[1] * ax[1]; r[1, 2] = t[1] * ax[2] + s * ax[0]
    r[2, 0] = t[2] * ax[0] + s * ax[1]; r[2, 1] = t[2] * ax[1] - s * ax[0]; r[2, 2] = c + t[2] * ax[2]
    out = np.empty((4, 4), dtype=F32)
    for k in range(3):
        out[k] = m[0] * r[k, 0] + m[1] * r[k, 1] + m[2] * r[k, 2]
    out[3] = m[3]
    return out


def translate(m: np.ndarray, v) -> np.ndarray:
    """glm::translate (ext/matrix_transform.inl:10-15)."""
    v = np.asarray(v, dtype=F32)
    out = m.astype(F32, copy=True)
    out[3] = m[0] * v[0] + m[1] * v[1] + m[2] * v[2] + m[3]
    return out


def mat_mul(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """glm a*b for [col,row] arrays."""
    out = np.empty((4, 4), dtype=F32)
    for c in range(4):
        out[c] = ((a[0] * b[c, 0] + a[1] * b[c, 1]) + a[2] * b[c, 2]) + a[3] * b[c, 3]
    return out


def mat_vec(m: np.ndarray, v) -> np.ndarray:
    v = np.asarray(v, dtype=F32)
    return ((m[0] * v[0] + m[1] * v[1]) + (m[2] * v[2] + m[3] * v[3])).astype(F32)


def bounds(pos: np.ndarray):
    return pos.min(axis=0).astype(F32), pos.max(axis=0).astype(F32)


def look_at_bounding_box(bb, up=(0.0, 1.0, 1.0), scale=1.0) -> np.ndarray:
    """``src/splat/camera.cpp:10-15``."""
    bmin, bmax = bb
    centre = (bmax + bmin) * F32(0.5)
    radius = np.sqrt(_dot3(bmax - bmin, bmax - bmin), dtype=F32) * F32(0.5)
    eye = centre - np.array([0, 0, F32(scale) * radius], dtype=F32)
    return look_at(eye, centre, up)


def fit_frustum_to_bounding_box(bb, fov_radians: float, aspect: float) -> np.ndarray:
    """``src/splat/geometry.cpp:9-24`` (near = R/tan(fov), far = near + 20R)."""
    bmin, bmax = bb
    d = (bmax - bmin).astype(F32)
    radius = np.sqrt(_dot3(d, d), dtype=F32) * F32(0.5)
    near = radius / _tanf(fov_radians)
    far = near + F32(20) * radius
    hw, hh = radius * F32(aspect), radius
    return frustum(-hw, hw, -hh, hh, near, far)


def mvp_start() -> np.ndarray:
    """The hard-coded initial view, ``src/main/splat.cpp:235-241``."""
    m = np.eye(4, dtype=F32)
    m[0, 0] = -1.0
    m[1, 1] = -0.09709989
    m[1, 2] = -0.99527466
    m[2, 1] = -0.99527466
    m[2, 2] = 0.09709989
    m[3, 2] = -5.1539507
    return m


def app_projection(g16: np.ndarray, fov_radians: float, width: int, height: int) -> np.ndarray:
    """Projection exactly as the app builds it (``src/main/splat.cpp:186-195``): look-at view of
    the bbox only serves to move the bbox corners to eye space for the frustum fit."""
    bb = bounds(g16[:, 0:3])
    model_view = look_at_bounding_box(bb, (0.0, 1.0, 1.0), 1.0)
    bb_cam = (mat_vec(model_view, [*bb[0], 1.0])[:3], mat_vec(model_view, [*bb[1], 1.0])[:3])
    return fit_frustum_to_bounding_box(bb_cam, fov_radians, width / float(height))


def orbit_view(k: int, steps: int = 32, base: np.ndarray | None = None) -> np.ndarray:
    """Config-2 orbit: ``glm::rotate(mvpStart, radians(360/steps * k), (0,1,0))``."""
    base = mvp_start() if base is None else base
    return rotate(base, np.radians(F32(360.0 / steps * k)), (0.0, 1.0, 0.0))


def ui_view(model_view, env_deg=0.0, env_deg2=0.0, X=640.0, Y=360.0, Z=1.0) -> np.ndarray:
    """``src/main/splat.cpp:312-314`` with the UI defaults of InterfaceServer.hpp:231-237."""
    v = mat_mul(model_view, rotate(np.eye(4, dtype=F32), np.radians(F32(env_deg)), (1.0, 0.0, 0.0)))
    v = rotate(v, np.radians(F32(env_deg2)), (0.0, 1.0, 0.0))
    return translate(v, (X / 50.0, Y / 50.0, -Z / 20.0 + 20.0))


DEFAULT_FOV = float(np.radians(F32(40.0)))   # src/main/splat.cpp:170,175


def multiview_cameras(g16: np.ndarray, n_orbit: int = 32, n_spiral: int = 32) -> list[np.ndarray]:
    """The multi-view batch of BASELINE configs[4] / SURVEY §8(d) config 5: `n_orbit` cameras on
    the app's orbit at 1.0x the bounding-box radius (``lookAtBoundingBox(bb,(0,1,1),1)`` rotated
    about y in equal steps, as ``src/main/splat.cpp:245-262`` turns the model), then `n_spiral`
    cameras on a two-turn spiral from 0.6x down to 0.05x the radius INSIDE the cloud, each looking
    at the centroid.  In-cloud views are the load-imbalance stress: Gaussians next to the eye
    cover many tiles and a few tiles hold most of the frame's list entries."""
    bb = bounds(g16[:, 0:3])
    bmin, bmax = bb
    centre = (bmax + bmin) * F32(0.5)
    radius = np.sqrt(_dot3(bmax - bmin, bmax - bmin), dtype=F32) * F32(0.5)
    base = look_at_bounding_box(bb, (0.0, 1.0, 1.0), 1.0)
    views = [rotate(base, np.radians(F32(360.0 / n_orbit * k)), (0.0, 1.0, 0.0)) for k in range(n_orbit)]
    for k in range(n_spiral):
        t = F32(k) / F32(max(n_spiral - 1, 1))
        r = radius * (F32(0.6) + (F32(0.05) - F32(0.6)) * t)
        ang = F32(4.0 * np.pi) * t
        eye = centre + r * np.array([np.cos(ang), F32(0.35) * np.sin(F32(0.5) * ang), np.sin(ang)], dtype=F32)
        views.append(look_at(eye.astype(F32), centre, (0.0, 1.0, 0.0)))
    return views


# ---- synthetic scenes (SURVEY §8d generator) ----------------------------------------------------
def synth_scene(n: int, seed: int) -> np.ndarray:
    """Seeded synthetic scene as (n,16) Gaussian3D records, from the HOST-ONLY generator library
    ``libgsplat_synth.so`` (csrc/synth.cpp compiled with g++, no CUDA).  Same stream of records as
    ``gsplat_synth_scene`` of the product library; this entry exists so that callers that must not
    map the product library (bench.py's CPU reference arm) can build the same scene."""
    import ctypes as C
    import os
    here = os.path.dirname(os.path.abspath(__file__))
    path = os.path.join(here, "libgsplat_synth.so")
    if not os.path.exists(path):
        raise FileNotFoundError(f"{path} is not built: run `make -C {os.path.join(here, 'csrc')}`")
    lib = C.CDLL(path)
    lib.gsplat_synth_scene.restype = C.c_int
    lib.gsplat_synth_scene.argtypes = [C.c_uint64, C.c_size_t, C.POINTER(C.c_float)]
    out = np.empty((n, 16), F32)
    rc = lib.gsplat_synth_scene(seed, n, out.ctypes.data_as(C.POINTER(C.c_float)))
    if rc:
        raise RuntimeError(f"gsplat_synth_scene failed: {rc}")
    return out


# ---- scene cache (same file as apps/splat_b200.cpp --cache) -----------------------
CACHE_MAGIC = b"GSPLB200"


def save_scene_cache(path: str, g16: np.ndarray) -> None:
    """The converted scene (16-float records after the PLY conversion and recentre + z-negate) as one
    flat little-endian file: magic | u32 version (1) | u32 0 | u64 n | f32 bbox min[3], max[3] | records."""
    g = np.ascontiguousarray(g16, dtype="<f4").reshape(-1, 16)
    bmin, bmax = bounds(g[:, 0:3]) if len(g) else (np.zeros(3, F32), np.zeros(3, F32))
    with open(path, "wb") as fh:
        fh.write(CACHE_MAGIC)
        fh.write(np.array([1, 0], dtype="<u4").tobytes())
        fh.write(np.array([len(g)], dtype="<u8").tobytes())
        fh.write(np.concatenate([bmin, bmax]).astype("<f4").tobytes())
        fh.write(g.tobytes())


def load_scene_cache(path: str) -> np.ndarray:
    """Records written by save_scene_cache / the C++ driver; raises ValueError on a foreign file."""
    with open(path, "rb") as fh:
        head = fh.read(48)
        if len(head) != 48 or head[:8] != CACHE_MAGIC or int(np.frombuffer(head, "<u4", 1, 8)[0]) != 1:
            raise ValueError(f"{path}: not a gsplat-b200 scene cache")
        n = int(np.frombuffer(head, "<u8", 1, 16)[0])
        g = np.fromfile(fh, dtype="<f4", count=n * 16)
    if g.size != n * 16:
        raise ValueError(f"{path}: truncated scene cache")
    return g.reshape(n, 16).astype(F32, copy=False)
