"""The host-side INPUTS of the path — PLY loading, the PLY -> Gaussian3D conversion, the cameras —
pinned to the reference's own code: happly through src/splat/file_io.cpp, lookAtBoundingBox
(src/splat/camera.cpp:10-15), fitFrustumToBoundingBox (src/splat/geometry.cpp:9-24) and glm, compiled
unchanged into oracle/_ref/libhost_ref.so (oracle/build_ref.sh).  scene.py (what the tests and
bench.py drive the library with) and apps/splat_b200 (the C++ host) must agree with it bit for bit
(matrices: to the ulp bound stated per test where numpy's and glibc's cos / sin / tan differ)."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ply_path
from oracle import pyoracle
from gaussian_splat_ipu_b200 import scene as sc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PLYS = ["point_cloud_9.ply", "point_cloud_10.ply", "point_cloud_11.ply", "point_cloud_12.ply"]


@pytest.fixture(scope="module")
def host():
    if not pyoracle.have_host_ref():
        pyoracle.build_ref()
    if not pyoracle.have_host_ref():
        pytest.skip("oracle/_ref/libhost_ref.so not built (needs /root/reference)")
    return pyoracle.HostRef()


def ulps(a, b):
    """Largest distance in float32 ulps between two arrays (0 = bit-identical; +0 / -0 count as equal)."""
    a, b = np.asarray(a, np.float32).ravel(), np.asarray(b, np.float32).ravel()
    ia, ib = a.view(np.int32).astype(np.int64), b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7FFFFFFF), ia)
    ib = np.where(ib < 0, -(ib & 0x7FFFFFFF), ib)
    return int(np.abs(ia - ib).max())


@pytest.mark.parametrize("name", PLYS)
def test_ply_loading_and_conversion_equal_happly_and_the_app(host, name):
    """scene.load_ply == the 14 properties happly returns for file_io.cpp:62-77; scene.gaussians_from_ply
    == the Gaussian3D records of src/main/splat.cpp:93-100,136-163 — every float bit for bit."""
    p = ply_path(name)
    if p is None:
        pytest.skip("PLY scenes not shipped")
    props, g16, bmin, bmax = host.load(p)
    ply = sc.load_ply(p)
    mine = np.stack([ply[k] for k in sc.PLY_PROPS], axis=1)
    assert mine.shape == props.shape and np.array_equal(mine.view(np.uint32), props.view(np.uint32))
    g = sc.gaussians_from_ply(ply)
    assert g.shape == g16.shape and np.array_equal(g.view(np.uint32), g16.view(np.uint32))
    b = sc.bounds(g[:, 0:3])
    assert np.array_equal(b[0], bmin) and np.array_equal(b[1], bmax)


@pytest.mark.parametrize("name", PLYS)
def test_cameras_equal_the_reference(host, name):
    """look_at_bounding_box, fit_frustum_to_bounding_box, the app's projection, the mvpStart orbit and
    the UI view against camera.cpp / geometry.cpp / glm.  Pure f32 arithmetic agrees bit for bit;
    matrices that go through tan / cos / sin may differ by the last bit of those functions (numpy's
    SIMD kernels vs glibc): bounded at 4 ulps per element and reported."""
    p = ply_path(name)
    if p is None:
        pytest.skip("PLY scenes not shipped")
    _, g16, bmin, bmax = host.load(p)
    bb = (bmin, bmax)
    assert ulps(sc.look_at_bounding_box(bb, (0.0, 1.0, 1.0), 1.0), host.look_at_bounding_box(bmin, bmax, (0, 1, 1), 1.0)) == 0
    worst = 0
    for W, H in ((1280, 720), (1920, 1080), (3840, 2160)):
        worst = max(worst, ulps(sc.fit_frustum_to_bounding_box(bb, sc.DEFAULT_FOV, W / float(H)),
                                host.fit_frustum(bmin, bmax, sc.DEFAULT_FOV, W / float(H))))
        worst = max(worst, ulps(sc.app_projection(g16, sc.DEFAULT_FOV, W, H), host.app_projection(bmin, bmax, sc.DEFAULT_FOV, W / float(H))))
    assert ulps(sc.mvp_start(), host.orbit_view(0.0)) == 0
    for k in range(32):
        worst = max(worst, ulps(sc.orbit_view(k, 32), host.orbit_view(np.float32(360.0 / 32 * k))))
    mv = sc.look_at_bounding_box(bb, (0.0, 1.0, 1.0), 1.0)
    for env, env2, X, Y, Z in ((0, 0, 640, 360, 1), (85.763603, 184.763657, 100, -40, 7), (-33.3, 12.5, 0, 0, 0)):
        worst = max(worst, ulps(sc.ui_view(mv, env, env2, X, Y, Z), host.ui_view(mv, env, env2, X, Y, Z)))
    base = sc.look_at_bounding_box(bb, (0.0, 1.0, 1.0), 1.0)
    for k in range(0, 64, 7):
        ang = np.radians(np.float32(360.0 / 64 * k))
        worst = max(worst, ulps(sc.rotate(base, ang, (0.0, 1.0, 0.0)), host.rotate(base, ang, (0, 1, 0))))
    print(f"{name}: worst camera-matrix difference {worst} ulp")
    assert worst == 0      # glibc host: scene.py calls the same tanf / cosf / sinf


def test_cpp_host_reads_the_same_scene_with_happly_and_with_its_own_reader(host, tmp_path):
    """apps/splat_b200 built WITH happly (the reference's loader, include/happly.h, when the reference
    tree is present at build time) and with its built-in reader write bit-identical scene caches, equal
    to the reference's records (no GPU needed: --cache-only stops before the device is touched)."""
    p = ply_path("point_cloud_11.ply")
    if p is None:
        pytest.skip("PLY scenes not shipped")
    r = subprocess.run(["make", "-C", os.path.join(ROOT, "apps")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    _, g16, _, _ = host.load(p)
    seen = 0
    for exe in ("splat_b200", "splat_b200_builtin_ply"):
        path = os.path.join(ROOT, "apps", exe)
        if not os.path.exists(path):
            continue
        cache = tmp_path / (exe + ".gsc")
        r = subprocess.run([path, "--input", p, "--cache", str(cache), "--cache-only"], capture_output=True, text=True, timeout=120)
        assert r.returncode == 0, r.stdout + r.stderr
        got = sc.load_scene_cache(str(cache))
        assert got.shape == g16.shape and np.array_equal(got.view(np.uint32), g16.view(np.uint32)), exe
        seen += 1
    assert seen >= 1
