"""CPU-only checks of the drop-in boundary: libgsplat_b200.so loads, exports every symbol that
include/gsplat_b200.h declares, the ctypes mirror binds exactly that set, and — with no CUDA
device — the product fails loudly instead of falling back to a CPU path."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from gaussian_splat_ipu_b200 import splatter as sp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "gsplat_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gsplat_[a-z0-9_]+)\s*\(", src)))


def have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_library_exports_every_declared_symbol():
    sp.build_library()
    lib = C.CDLL(sp.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in gsplat_b200.h but not exported"


def test_ctypes_mirror_binds_the_declared_set():
    assert sorted(sp.SYMBOLS) == declared_symbols()
    lib = sp.load_library()
    assert lib.gsplat_abi_version() == 4


def test_config_defaults_are_the_reference_constants():
    lib = sp.load_library()
    cfg = sp.Config()
    assert lib.gsplat_config_default(C.byref(cfg)) == 0
    # tile_config.hpp:5-15 (1280x720, 40x36 tiles of 32x20), codelets.cpp:622 (clipSize 15)
    assert (cfg.width, cfg.height, cfg.tile_w, cfg.tile_h) == (1280, 720, 32, 20)
    assert cfg.guard_factor == 15.0 and cfg.device == 0
    assert (cfg.band_row_begin, cfg.band_row_end) == (0, -1)        # -1: all tile rows; (0, 0) is an empty band
    assert lib.gsplat_config_default(None) == -1


def test_header_cites_the_reference_interface():
    src = open(HEADER).read()
    for cite in ("ipu_rasteriser.hpp:20-55", "ipu_rasteriser.cpp:86-110", "ipu_rasteriser.cpp:115-144",
                 "codelets.cpp:143-641", "tile_config.hpp"):
        assert cite in src


def test_synth_scene_is_deterministic_and_in_spec():
    """Host-only helper (SURVEY §8d generator): seeded, reproducible, reference record layout."""
    a = sp.synth_scene(5000, 0xB2000003)
    b = sp.synth_scene(5000, 0xB2000003)
    c = sp.synth_scene(5000, 0xB2000004)
    assert np.array_equal(a, b) and not np.array_equal(a, c)
    assert a.shape == (5000, 16) and np.isfinite(a).all()
    assert (a[:, 3] == 1.0).all()                                   # mean.w
    assert np.array_equal(a[:, 15], np.arange(1, 5001, dtype=np.float32))   # gid = i+1
    assert (a[:, 4:7] >= 0).all() and (a[:, 4:7] < 1).all()        # rgb ~ U[0,1)
    assert (a[:, 7] >= 0.05).all() and (a[:, 7] < 1).all()         # opacity ~ U[0.05,1)
    assert (a[:, 12:15] >= -13).all() and (a[:, 12:15] <= -0.5).all()
    assert np.allclose(np.linalg.norm(a[:, 8:12], axis=1), 1.0, atol=1e-5)
    # recentred on the bbox centroid (src/main/splat.cpp:93-100)
    assert np.allclose(a[:, :3].max(0) + a[:, :3].min(0), 0, atol=1e-5)
    # a prefix of a longer scene is the same stream
    assert np.array_equal(sp.synth_scene(100, 7)[:, 4:15], sp.synth_scene(200, 7)[:100, 4:15])


def test_host_only_synth_library_is_the_same_generator():
    """libgsplat_synth.so (csrc/synth.cpp built with g++, no CUDA: what bench.py's CPU reference arm
    loads so that it maps no product library) emits bit-identical records to gsplat_synth_scene of
    the product library."""
    from gaussian_splat_ipu_b200 import scene as sc
    for n, seed in ((1, 1), (4097, 0xB2000006), (50000, 7)):
        a, b = sc.synth_scene(n, seed), sp.synth_scene(n, seed)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))


def test_reference_arm_does_not_map_the_product_library():
    """bench.py --impl reference builds its scene without dlopen-ing libgsplat_b200.so."""
    import subprocess
    import sys
    code = ("import sys; sys.argv=['bench.py','--impl','reference','--workload','synthetic-200k-720p','--steps','2','--warmup','1'];"
            "import runpy\n"
            "try:\n runpy.run_path('bench.py', run_name='__main__')\nexcept SystemExit: pass\n"
            "maps=open('/proc/self/maps').read()\n"
            "sys.stderr.write('MAPPED' if 'libgsplat_b200' in maps else 'CLEAN')")
    r = subprocess.run([sys.executable, "-c", code], cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stderr.strip().endswith("CLEAN"), r.stderr[-2000:]


@pytest.mark.skipif(have_gpu(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback_without_a_device():
    g = sp.synth_scene(10, 1)
    with pytest.raises(sp.GsplatError) as e:
        sp.B200Splatter(g, 1280, 720)
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)


def test_argument_errors_do_not_need_a_device():
    lib = sp.load_library()
    cfg = sp.Config()
    lib.gsplat_config_default(C.byref(cfg))
    h = C.c_void_p()
    cfg.width = 100                                   # 32 does not divide 100
    assert lib.gsplat_create(C.byref(cfg), C.byref(h)) == -1 and not h.value
    assert b"tile size" in lib.gsplat_last_error(None)
    assert lib.gsplat_create(None, C.byref(h)) == -1
    assert lib.gsplat_render(None) == -1 and lib.gsplat_sync(None) == -1


def test_product_does_not_import_the_oracle():
    """The oracle is test infrastructure: nothing under the package may import or load it."""
    pkg = os.path.join(ROOT, "gaussian_splat_ipu_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".hpp", ".inc", "Makefile")):
                txt = open(os.path.join(dp, f), errors="replace").read()
                assert "pyoracle" not in txt and "liboracle" not in txt and "splat_ref" not in txt, os.path.join(dp, f)


def _build_cpp(tmpdir, name):
    import subprocess
    exe = os.path.join(tmpdir, name)
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    libdir = os.path.join(ROOT, "gaussian_splat_ipu_b200")
    cmd = [cxx, "-std=c++17", "-O1", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cpp", name + ".cpp"), "-o", exe, "-L", libdir, "-lgsplat_b200",
           f"-Wl,-rpath,{libdir}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def _build_shim_smoke(tmpdir):
    return _build_cpp(tmpdir, "shim_smoke")


def test_group_host_compiles_and_group_arguments_are_checked(tmp_path):
    """tests/cpp/group_smoke.cpp (a C++ host over gsplat_create_group) compiles warning-free and
    links; bad group arguments are refused without touching a device."""
    _build_cpp(str(tmp_path), "group_smoke")
    lib = sp.load_library()
    cfg = sp.Config()
    lib.gsplat_config_default(C.byref(cfg))
    h = C.c_void_p()
    one = (C.c_int * 2)(0, 0)
    assert lib.gsplat_create_group(C.byref(cfg), one, 2, 1, C.byref(h)) == -1 and b"twice" in lib.gsplat_group_last_error(None)
    assert lib.gsplat_create_group(C.byref(cfg), one, 0, 1, C.byref(h)) == -1
    assert lib.gsplat_create_group(C.byref(cfg), one, 1, 7, C.byref(h)) == -1
    assert lib.gsplat_group_render_streamed(None, None) == -1 and lib.gsplat_group_size(None) == 0


def test_static_sort_argument_errors_do_not_need_a_device():
    """gsplat_sort_pairs_static checks its arguments before it touches a device: sizes beyond the static
    scratch array, bit ranges, pointers."""
    err = C.create_string_buffer(200)
    f = sp.load_library().gsplat_sort_pairs_static
    assert f(0, 256, 512, 768, 1024, sp.SORT_STATIC_MAX + 1, 0, 32, None, err, len(err)) == -3
    assert b"GSPLAT_SORT_STATIC_MAX" in err.value and str(sp.SORT_STATIC_MAX).encode() in err.value
    assert f(0, 256, 512, 768, 1024, 8, 4, 4, None, err, len(err)) == -1 and b"bit range" in err.value
    assert f(0, 256, 512, 768, 1024, 8, 0, 33, None, err, len(err)) == -1
    assert f(0, 0, 512, 768, 1024, 8, 0, 32, None, err, len(err)) == -1 and b"null" in err.value
    assert f(0, 260, 512, 768, 1024, 8, 0, 32, None, err, len(err)) == -1 and b"aligned" in err.value
    assert f(0, 256, 512, 256, 1024, 8, 0, 32, None, err, len(err)) == -1 and b"distinct" in err.value
    assert f(0, 0, 0, 0, 0, 0, 0, 32, None, None, 0) == 0          # nothing to sort
    assert f(0, 256, 512, 768, 1024, sp.SORT_STATIC_MAX + 1, 0, 32, None, None, 0) == -3   # no message buffer


@pytest.mark.skipif(have_gpu(), reason="checks the no-GPU failure mode")
def test_cpp_shim_compiles_and_fails_loudly_without_device(tmp_path):
    """include/gsplat_b200.hpp (the IpuSplatter-shaped C++ host mirror) compiles warning-free,
    links against the library, and without a device throws instead of falling back."""
    import subprocess
    sp.build_library()
    exe = _build_shim_smoke(str(tmp_path))
    r = subprocess.run([exe, "1000"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 3 and "no CPU fallback" in r.stdout, (r.returncode, r.stdout, r.stderr)


def test_cpp_host_driver_builds_and_fails_loudly_without_a_device():
    """apps/splat_b200 links against the C ABI; with no CUDA device it reports the library's error
    (exit code 3) instead of rendering on the CPU."""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sp.build_library()
    subprocess.run(["make", "-C", os.path.join(root, "apps")], check=True, capture_output=True)
    r = subprocess.run([os.path.join(root, "apps", "splat_b200"), "--synthetic", "1000", "--frames", "2"],
                       capture_output=True, text=True, timeout=60)
    import ctypes
    try:
        ctypes.CDLL("libcuda.so.1")
        have_driver = True
    except OSError:
        have_driver = False
    if r.returncode == 0:
        assert r.stdout.startswith("ok ")          # a GPU box
    else:
        assert r.returncode == 3 and r.stdout.startswith("no-device:"), r.stdout + r.stderr
    assert subprocess.run([os.path.join(root, "apps", "splat_b200")], capture_output=True).returncode == 2   # usage


def test_scene_cache_round_trip(tmp_path):
    """the converted scene as a flat file (shared with apps/splat_b200 --cache)."""
    from gaussian_splat_ipu_b200 import scene as sc
    from gaussian_splat_ipu_b200.splatter import synth_scene
    g = synth_scene(777, 9)
    p = str(tmp_path / "scene.gsc")
    sc.save_scene_cache(p, g)
    assert os.path.getsize(p) == 48 + g.size * 4
    h = sc.load_scene_cache(p)
    assert h.dtype == np.float32 and np.array_equal(h.view(np.uint32), g.view(np.uint32))
    with open(p, "r+b") as fh:
        fh.write(b"NOTACACHE")
    with pytest.raises(ValueError):
        sc.load_scene_cache(p)
    sc.save_scene_cache(p, g)
    with open(p, "r+b") as fh:
        fh.truncate(48 + 100)
    with pytest.raises(ValueError):
        sc.load_scene_cache(p)
