"""The product's ORBmatcher against the UNMODIFIED reference src/ORBmatcher.cc (oracle/_ref/libmatcher_ref.so: the reference's file
compiled where it lies over the cv shim and stand-in Frame / KeyFrame / MapPoint objects, oracle/ref_shim/slam) on the same objects.
GPU leg: tests/native/reference_matcher_test.cpp runs every search both ways and exits non-zero on any difference.
CPU leg: the same program linked against tests/native/orbx_on_oracle.cpp -- the 21 C-ABI entry points it needs, implemented by the CPU
oracle (test infrastructure, never shipped) -- instead of liborbx.so: the headers' flattening code and the oracle's searches against the
reference's own matcher, without a GPU.  With the GPU suite's CUDA == oracle tests this closes the triangle."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
REFLIB = os.path.join(ROOT, "oracle", "_ref", "libmatcher_ref.so")
# The test programs are compiled the way the reference's tree is (CMakeLists.txt:26-27: -O3 -march=native; x86-64-v3 here as for oracle/_ref): the
# drop-in templates hold the reference's float expressions (fx*xc*invzc+cx, u-mbf*invzc ...), and whether g++ contracts them into fma must be the
# same on both sides, as it is when a maintainer compiles the swapped header inside the reference's own build.  (Found by
# tools/fuzz_reference_objects.py: at -O1 without FMA, 5 of 80 random scenes differed in one or two window-boundary cases.)
REF_FLAGS = ["-O3", "-march=x86-64-v3"]


def _reflib():
    """oracle/_ref/libmatcher_ref.so: built here from /root/reference when that exists, else the prebuilt file that travelled."""
    if os.path.isdir("/root/reference"):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "_ref"], stdout=subprocess.DEVNULL)
    if not os.path.exists(REFLIB):
        pytest.skip("oracle/_ref/libmatcher_ref.so is not built and /root/reference is absent")
    return REFLIB


def _build(tmp_path, on_oracle=False):
    import orbx
    _reflib()
    if not os.path.exists(orbx.LIB_PATH):
        sys.path.insert(0, ROOT)
        import __graft_entry__ as g
        g.build()
    exe = str(tmp_path / ("refmatch_cpu" if on_oracle else "refmatch"))
    libdir, refdir, oradir = os.path.dirname(orbx.LIB_PATH), os.path.dirname(REFLIB), os.path.join(ROOT, "oracle")
    if on_oracle:       # the C ABI the headers call, served by the CPU oracle
        import oracle_py  # noqa: builds oracle/liborb_oracle.so when it is missing
        subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-fPIC", "-shared", "-I", os.path.join(ROOT, "include"), "-I", oradir,
                               "-o", str(tmp_path / "liborbx_on_oracle.so"), os.path.join(ROOT, "tests", "native", "orbx_on_oracle.cpp"),
                               "-L", oradir, "-lorb_oracle", "-Wl,-rpath," + oradir])
        link = ["-L", str(tmp_path), "-lorbx_on_oracle", "-Wl,-rpath," + str(tmp_path)]
    else:
        link = ["-L", libdir, "-lorbx", "-Wl,-rpath," + libdir]
    subprocess.check_call(["g++", "-std=c++11"] + REF_FLAGS + ["-Wall", "-Werror", "-DORBX_USE_OPENCV", "-I", os.path.join(ROOT, "include"),
                           "-I", os.path.join(ROOT, "oracle", "ref_shim"), "-I", os.path.join(ROOT, "oracle", "ref_shim", "slam"),
                           "-o", exe, os.path.join(ROOT, "tests", "native", "reference_matcher_test.cpp")] + link +
                          ["-L", refdir, "-lmatcher_ref", "-Wl,-rpath," + refdir])
    return exe


def test_reference_matcher_library_exports_every_entry_point(tmp_path):
    lib = _reflib()
    syms = subprocess.run(["nm", "-DC", lib], capture_output=True, text=True, check=True).stdout
    for name in ("refm::SearchByProjection(float, bool, ORB_SLAM2_REF::Frame&, std::vector", "refm::SearchByProjection(float, bool, ORB_SLAM2_REF::Frame&, ORB_SLAM2_REF::Frame const&",
                 "refm::SearchByProjection(float, bool, ORB_SLAM2_REF::Frame&, ORB_SLAM2_REF::KeyFrame*", "refm::SearchByProjection(float, bool, ORB_SLAM2_REF::KeyFrame*, cv::Mat",
                 "refm::SearchByBoW(float, bool, ORB_SLAM2_REF::KeyFrame*, ORB_SLAM2_REF::Frame&", "refm::SearchByBoW(float, bool, ORB_SLAM2_REF::KeyFrame*, ORB_SLAM2_REF::KeyFrame*",
                 "refm::SearchForInitialization", "refm::SearchForTriangulation", "refm::SearchBySim3", "refm::Fuse(float, bool, ORB_SLAM2_REF::KeyFrame*, std::vector",
                 "refm::Fuse(float, bool, ORB_SLAM2_REF::KeyFrame*, cv::Mat", "refm::DescriptorDistance",
                 # the reference's own class, under the renamed namespace
                 "ORB_SLAM2_REF::ORBmatcher::SearchByBoW", "ORB_SLAM2_REF::ORBmatcher::ComputeThreeMaxima", "ORB_SLAM2_REF::ORBmatcher::DescriptorDistance"):
        assert name in syms, name
    _build(tmp_path)        # the test program compiles and links against it


def _views(tmp_path, w, h, seed, dx, dy, kind="rects"):
    from orbx import synth
    a = getattr(synth, kind)(w, h, seed)
    rng = np.random.default_rng(seed)
    b = np.clip(np.roll(a, (dy, dx), axis=(0, 1)).astype(np.int16) + rng.integers(-3, 4, a.shape), 0, 255).astype(np.uint8)
    (tmp_path / "a.raw").write_bytes(a.tobytes())
    (tmp_path / "b.raw").write_bytes(b.tobytes())


SCENES = [(640, 480, 1000, 2100, 3, 1), (752, 480, 1200, 2200, -4, 2), (640, 480, 2000, 2300, 2, -3)]


@pytest.fixture(scope="module")
def cpu_exe(tmp_path_factory):
    return _build(tmp_path_factory.mktemp("refmatch_cpu"), on_oracle=True)


@pytest.mark.parametrize("w,h,nf,seed,dx,dy,kind", [s + ("rects",) for s in SCENES] +
                         [(400, 300, 500, 2400, 5, 4, "rects"), (640, 480, 1000, 2500, 2, 2, "texture"), (512, 384, 800, 2600, -3, -2, "texture"),
                          (752, 480, 1500, 2700, 6, 0, "rects"), (640, 480, 3000, 2800, 1, 1, "texture")])
def test_oracle_searches_equal_the_reference_matcher_on_cpu(tmp_path, cpu_exe, w, h, nf, seed, dx, dy, kind):
    """All eleven searches: the drop-in headers over the CPU oracle against the unmodified reference ORBmatcher.cc, same objects."""
    _views(tmp_path, w, h, seed, dx, dy, kind)
    r = subprocess.run([cpu_exe, str(tmp_path / "a.raw"), str(tmp_path / "b.raw"), str(w), str(h), str(nf), str(dx), str(dy)], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.rstrip().endswith("ok"), (r.returncode, r.stdout[-3000:], r.stderr[-3000:])
    assert r.stdout.count("SearchBySim3") == 4 and r.stdout.count("Fuse(KF, MPs)") == 2


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,nf,seed,dx,dy", SCENES)
def test_searches_equal_the_reference_matcher(tmp_path, w, h, nf, seed, dx, dy):
    exe = _build(tmp_path)
    _views(tmp_path, w, h, seed, dx, dy)
    r = subprocess.run([exe, str(tmp_path / "a.raw"), str(tmp_path / "b.raw"), str(w), str(h), str(nf), str(dx), str(dy)], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.rstrip().endswith("ok"), (r.returncode, r.stdout[-3000:], r.stderr[-3000:])


# ---- the same comparison on the reference's REAL objects: tests/native/reference_objects_test.cpp compiles the product's drop-in ORBmatcher
# against the reference's own Frame.h / KeyFrame.h / MapPoint.h / ORBmatcher.h and links the reference's own object code
# (oracle/_ref/libslam_ref.so: unmodified Frame.cc, KeyFrame.cc, MapPoint.cc, ORBmatcher.cc, ORBextractor.cc, DBoW2).  Compiling it needs the
# reference tree, so the GPU leg runs the executable `make -C oracle _ref_tests` built where that tree exists.
REAL_EXE = os.path.join(ROOT, "oracle", "_ref", "reference_objects_test")
REF_TREE = os.environ.get("ORB_REFERENCE_ROOT", "/root/reference")


@pytest.fixture(scope="module")
def real_cpu_exe(tmp_path_factory):
    if not os.path.isdir(os.path.join(REF_TREE, "include")):
        pytest.skip("the reference tree is not here: its headers are needed to compile this test")
    tmp = tmp_path_factory.mktemp("refobjects_cpu")
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "_ref/libslam_ref.so"], stdout=subprocess.DEVNULL)
    import oracle_py  # noqa: builds oracle/liborb_oracle.so when it is missing
    oradir, shim = os.path.join(ROOT, "oracle"), os.path.join(ROOT, "oracle", "ref_shim")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-fPIC", "-shared", "-I", os.path.join(ROOT, "include"), "-I", oradir,
                           "-o", str(tmp / "liborbx_on_oracle.so"), os.path.join(ROOT, "tests", "native", "orbx_on_oracle.cpp"),
                           "-L", oradir, "-lorb_oracle", "-Wl,-rpath," + oradir])
    exe = str(tmp / "refobjects_cpu")
    subprocess.check_call(["g++", "-std=c++11"] + REF_FLAGS + ["-Wall", "-DORBX_USE_OPENCV", "-Wno-overloaded-virtual", "-I", os.path.join(shim, "slam_full"), "-I", shim,
                           "-I", os.path.join(REF_TREE, "include"), "-I", REF_TREE, "-o", exe, os.path.join(ROOT, "tests", "native", "reference_objects_test.cpp"),
                           "-L", str(tmp), "-lorbx_on_oracle", "-Wl,-rpath," + str(tmp), "-L", os.path.join(oradir, "_ref"), "-lslam_ref",
                           "-Wl,-rpath," + os.path.join(oradir, "_ref")])
    return exe


def _run_real(exe, tmp_path, w, h, nf, seed, dx, dy, kind="rects"):
    _views(tmp_path, w, h, seed, dx, dy, kind)
    r = subprocess.run([exe, str(tmp_path / "a.raw"), str(tmp_path / "b.raw"), str(w), str(h), str(nf), str(dx), str(dy)], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.rstrip().endswith("ok"), (r.returncode, r.stdout[-3000:], r.stderr[-3000:])
    assert r.stdout.count("SearchBySim3") == 4 and r.stdout.count("Fuse(KF, MPs)") == 2 and r.stdout.count("SearchForTriangulation") == 4


@pytest.mark.parametrize("w,h,nf,seed,dx,dy,kind", [s + ("rects",) for s in SCENES] +
                         [(400, 300, 500, 2400, 5, 4, "rects"), (640, 480, 1000, 2500, 2, 2, "texture"), (752, 480, 1500, 2700, 6, 0, "rects")])
def test_drop_in_headers_on_the_reference_s_real_objects_cpu(tmp_path, real_cpu_exe, w, h, nf, seed, dx, dy, kind):
    """All eleven searches, the product's ORBmatcher (over the CPU oracle's C ABI) and the reference's, same call lines, on real
    ORB_SLAM2 Frame / KeyFrame / MapPoint objects."""
    _run_real(real_cpu_exe, tmp_path, w, h, nf, seed, dx, dy, kind)


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,nf,seed,dx,dy", SCENES)
def test_drop_in_headers_on_the_reference_s_real_objects(tmp_path, w, h, nf, seed, dx, dy):
    if os.path.isdir(os.path.join(REF_TREE, "include")):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "_ref_tests"], stdout=subprocess.DEVNULL)
    if not os.path.exists(REAL_EXE):
        pytest.skip("oracle/_ref/reference_objects_test was not built (make -C oracle _ref_tests needs the reference tree)")
    _run_real(REAL_EXE, tmp_path, w, h, nf, seed, dx, dy)


@pytest.mark.gpu
def test_drop_in_headers_on_real_objects_random_scenes():
    """tools/fuzz_reference_objects.py with a fixed seed: six random scenes (kinds, sizes, feature counts 300-3000, shifts) through the CUDA
    searches, on the reference's real objects (profiles/r02_fuzz_reference_objects_gpu.log holds this very run)."""
    if not os.path.exists(REAL_EXE):
        pytest.skip("oracle/_ref/reference_objects_test was not built (make -C oracle _ref_tests needs the reference tree)")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_reference_objects.py"), REAL_EXE, "6", "31"], capture_output=True, text=True)
    assert r.returncode == 0 and "6 runs, 0 with a difference" in r.stdout, (r.stdout[-2000:], r.stderr[-2000:])
