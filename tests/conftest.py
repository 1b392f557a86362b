"""Shared fixtures.  `gpu` marks tests that need a B200 (run with `-m gpu` on the GPU box);
everything else runs on CPU.  The oracle (oracle/) is the checker, never the thing under test's
implementation."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import pyoracle  # noqa: E402
from gaussian_splat_ipu_b200 import scene as sc  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def port():
    return pyoracle.Oracle("port")


@pytest.fixture(scope="session")
def ref():
    if not pyoracle.have_ref():
        pyoracle.build_ref()
    if not pyoracle.have_ref():
        pytest.skip("oracle/_ref/libsplat_ref.so not built (needs /root/reference)")
    return pyoracle.Oracle("reference")


def ply_path(name):
    p = os.path.join(pyoracle.REF_DATA, name)
    return p if os.path.exists(p) else None


def ply_scene(name="point_cloud_10.ply"):
    """(g16, view, proj) for a shipped PLY with the app's default camera, or None if absent."""
    p = ply_path(name)
    if p is None:
        return None
    g = sc.gaussians_from_ply(sc.load_ply(p))
    return g, sc.mvp_start(), sc.app_projection(g, sc.DEFAULT_FOV, 1280, 720)


def synthetic_camera(g16, width, height, fov=sc.DEFAULT_FOV):
    """Camera of SURVEY §8d: lookAtBoundingBox(bb,(0,1,1),1) + fitFrustumToBoundingBox."""
    bb = sc.bounds(g16[:, 0:3])
    view = sc.look_at_bounding_box(bb, (0.0, 1.0, 1.0), 1.0)
    proj = sc.app_projection(g16, fov, width, height)
    return view, proj


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name))


def compare_lists(off_a, lst_a, off_b, lst_b):
    assert np.array_equal(off_a, off_b), "per-tile counts differ"
    assert np.array_equal(lst_a, lst_b), "per-tile lists differ"
