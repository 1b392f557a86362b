"""frames streamed to `remote_render_ui` over packetcomms.  apps/stream_server.hpp
(B200StreamServer) keeps the reference's InterfaceServer surface and wire protocol
(include/remote_ui/InterfaceServer.hpp:24-43 packet types, :110-203 state ingestion, :322-332 preview +
histogram, external/packetcomms/src/Sync.cpp `ready` handshake) on the reference's own transport,
compiled unchanged from the reference tree.  The client side of these tests (tests/cpp/ui_client.cpp)
is written with packetcomms' own PacketMuxer / PacketDemuxer / TcpSocket, as remote_render_ui is."""
import os
import socket
import subprocess
import time

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
APPS = os.path.join(ROOT, "apps")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _programs():
    if not os.path.exists(os.path.join(APPS, "ui_client")):
        subprocess.run(["make", "-C", APPS], capture_output=True)
    if not os.path.exists(os.path.join(APPS, "ui_client")):
        pytest.skip("packetcomms programs not built (need the reference tree at build time)")


def test_wire_protocol_loopback(tmp_path):
    """No GPU: a stand-in render loop behind B200StreamServer paints the UI state it received into the
    frame it sends.  The client does the `ready` handshake, sends every float state packet the viewer's
    sliders send (fov in degrees), and gets back the histogram (cereal vector<uint32>) and preview
    frames carrying exactly that state; `stop` ends the server's loop."""
    _programs()
    port = _free_port()
    W, H = 96, 60
    srv = subprocess.Popen([os.path.join(APPS, "stream_loopback"), str(port), str(W), str(H)], stdout=subprocess.PIPE,
                           stderr=subprocess.PIPE, text=True)
    try:
        out_img, out_hist = tmp_path / "f.bgr", tmp_path / "h.u32"
        r = subprocess.run([os.path.join(APPS, "ui_client"), str(port), "5", str(out_img), str(out_hist), "X=100", "Y=50",
                            "Z=3", "lambda1=2.5", "lambda2=0.7", "gamma=1.8", "env_rotation=30", "env_rotation_2=45",
                            "fov=60", "exposure=1.5"], capture_output=True, text=True, timeout=60)
        line = [l for l in r.stdout.splitlines() if l.startswith(("ok", "incomplete"))][-1].split()
        assert r.returncode == 0 and line[0] == "ok", r.stdout[-1500:] + r.stderr[-500:]
        got = {line[i]: line[i + 1] for i in range(1, len(line) - 1, 2)}
        assert (int(got["width"]), int(got["height"]), int(got["format"])) == (W, H, 0)
        assert abs(float(got["fov_echo"]) - 0.6981317) < 1e-6          # the server's initial fov, in radians
        img = np.fromfile(out_img, np.uint8)
        assert img.size == W * H * 3
        assert list(img[:9]) == [100, 50, 3, 25, 7, 18, 30, 45, 60]    # X Y Z | lambda1 lambda2 gamma (x10) | env, env2, fov deg
        hist = np.fromfile(out_hist, np.uint32)
        assert np.array_equal(hist, np.arange(hist.size, dtype=np.uint32) * 3 + 1) and hist.size == (W // 32) * (H // 20)
        so, _ = srv.communicate(timeout=20)
        assert srv.returncode == 0 and "stop 1" in so, so
    finally:
        if srv.poll() is None:
            srv.kill()


@pytest.mark.gpu
def test_interactive_driver_streams_the_oracle_frame(tmp_path):
    """apps/splat_b200 --ui-port: the interactive loop of src/main/splat.cpp:245-322 on the B200 core.  A
    packetcomms client sets the camera sliders, fov and lambda1 as the viewer does and receives frames;
    the frame for the final state is byte for byte the a10 conversion of the ORACLE's framebuffer for
    the UI camera (splat.cpp:312-314), and the histogram is the oracle's per-tile list lengths."""
    from oracle import pyoracle
    from gaussian_splat_ipu_b200 import scene as sc
    _programs()
    ply = os.path.join(pyoracle.REF_DATA, "point_cloud_10.ply")
    if not os.path.exists(ply):
        pytest.skip("PLY scenes not shipped")
    port = _free_port()
    W, H = 1280, 720
    env, env2, X, Y, Z, lam, fov_deg = 95.0, 184.5, 20.0, -35.0, 300.0, 10.0, 40.0
    srv = subprocess.Popen([os.path.join(APPS, "splat_b200"), "--input", ply, "--ui-port", str(port)], stdout=subprocess.PIPE,
                           stderr=subprocess.PIPE, text=True)
    try:
        out_img, out_hist = tmp_path / "f.bgr", tmp_path / "h.u32"
        r = subprocess.run([os.path.join(APPS, "ui_client"), str(port), "6", str(out_img), str(out_hist), f"env_rotation={env}",
                            f"env_rotation_2={env2}", f"X={X}", f"Y={Y}", f"Z={Z}", f"lambda1={lam}", f"fov={fov_deg}"],
                           capture_output=True, text=True, timeout=120)
        assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-500:]
        so, se = srv.communicate(timeout=30)
        assert srv.returncode == 0 and so.strip().startswith("ok ui frames"), so + se[-800:]
    finally:
        if srv.poll() is None:
            srv.kill()
    g = sc.gaussians_from_ply(sc.load_ply(ply))
    bb = sc.bounds(g[:, 0:3])
    model_view = sc.look_at_bounding_box(bb, (0.0, 1.0, 1.0), 1.0)
    # InterfaceServer.hpp:189-196: state.fov = state.fov * (M_PI / 180.f) — a double product rounded once to float
    fov = np.float32(np.float64(np.float32(fov_deg)) * (np.pi / np.float64(np.float32(180.0))))
    view = sc.ui_view(model_view, env, env2, X, Y, Z)
    proj = sc.app_projection(g, float(fov), W, H)
    f = pyoracle.Oracle("port").frame(g, view, proj, float(fov), float(np.float32(lam) / np.float32(10.0)), W, H)
    img = np.fromfile(out_img, np.uint8).reshape(H, W, 3)
    want = pyoracle.Oracle("port").to_bgr8(f.fb, W, H)
    assert int(f.drawn.sum()) > 1000 and want.any(), "the UI camera must see the scene"
    assert np.array_equal(np.fromfile(out_hist, np.uint32), np.diff(f.offsets).astype(np.uint32))
    assert np.array_equal(img, want), f"{int((img != want).sum())} bytes differ from the oracle's frame"
