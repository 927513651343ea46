"""The drop-in C++ headers (include/ORBextractor.h, include/ORBmatcher.h) used the way Frame.cc / Tracking.cc use the
reference classes (tests/native/dropin_test.cpp).  CPU leg: the program compiles with g++ alone, links against
liborbx.so, and -- there being no CPU fallback -- fails loudly without a device.  GPU leg: its outputs (keypoints,
descriptors, mvImagePyramid, a brute-force self match, DescriptorDistance) equal the oracle's."""
import os
import struct
import subprocess

import numpy as np
import pytest

import oracle_py as O
import orbx
from orbx import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(tmp_path):
    if not os.path.exists(orbx.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    exe = str(tmp_path / "dropin_test")
    libdir = os.path.dirname(orbx.LIB_PATH)
    subprocess.check_call(["g++", "-std=c++14", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"), "-o", exe,
                           os.path.join(ROOT, "tests", "native", "dropin_test.cpp"), "-L", libdir, "-lorbx", "-Wl,-rpath," + libdir])
    return exe


def _build_native(tmp_path, source, name, opencv_shaped=False):
    """g++ one of tests/native/*.cpp against include/.  opencv_shaped: compile the ORBX_USE_OPENCV branch of the headers (what a
    user of the reference builds) against an OpenCV-shaped core.hpp -- CV_8U as a macro, cv::_InputArray / _OutputArray classes,
    MatStep -- for which oracle/ref_shim serves (this image has no OpenCV C++ headers)."""
    if not os.path.exists(orbx.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    exe = str(tmp_path / name)
    libdir = os.path.dirname(orbx.LIB_PATH)
    cmd = ["g++", "-std=c++11", "-O1", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "tests", "native")]
    if opencv_shaped:
        cmd += ["-DORBX_USE_OPENCV", "-I", os.path.join(ROOT, "oracle", "ref_shim")]
    subprocess.check_call(cmd + ["-o", exe, os.path.join(ROOT, "tests", "native", source), "-L", libdir, "-lorbx", "-Wl,-rpath," + libdir])
    return exe


@pytest.mark.parametrize("opencv_shaped", [False, True], ids=["bundled_shim", "opencv_shaped_headers"])
@pytest.mark.parametrize("source", ["dropin_test.cpp", "reference_api_test.cpp"])
def test_headers_compile_in_both_modes(tmp_path, source, opencv_shaped):
    """include/ORBextractor.h and include/ORBmatcher.h -- with the reference's call lines of Tracking.cc:742-743, :910, :1028,
    :1336 and LoopClosing.cc:275 written verbatim in reference_api_test.cpp -- compile and link with the bundled cv shim and in
    the ORBX_USE_OPENCV branch."""
    _build_native(tmp_path, source, "t_" + source[:-4], opencv_shaped)


def _has_gpu():
    import torch
    return torch.cuda.is_available()


def test_dropin_compiles_links_and_has_no_cpu_fallback(tmp_path):
    exe = _build(tmp_path)
    if _has_gpu():
        pytest.skip("GPU present: covered by the gpu leg")
    img = synth.rects(320, 240, 7)
    raw = tmp_path / "in.raw"
    raw.write_bytes(img.tobytes())
    r = subprocess.run([exe, str(raw), "320", "240", "500", str(tmp_path / "out.bin")], capture_output=True, text=True)
    assert r.returncode == 1 and "orbx_create" in r.stderr, (r.returncode, r.stderr)


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,nf,seed", [(640, 480, 1000, 1000), (752, 480, 1200, 2000)])
def test_dropin_matches_oracle(tmp_path, w, h, nf, seed):
    exe = _build(tmp_path)
    img = synth.rects(w, h, seed)
    raw = tmp_path / "in.raw"
    raw.write_bytes(img.tobytes())
    out = tmp_path / "out.bin"
    import scenes
    voc = scenes.synthetic_vocabulary(seed, 10, 5)
    scenes.write_vocabulary_text(str(tmp_path / "voc.txt"), *voc)
    subprocess.check_call([exe, str(raw), str(w), str(h), str(nf), str(out), str(tmp_path / "voc.txt")])
    b = out.read_bytes()
    n, = struct.unpack_from("<i", b, 0)
    p = 4
    kps = np.frombuffer(b, dtype=orbx.KP_DTYPE, count=n, offset=p); p += 28 * n
    desc = np.frombuffer(b, dtype=np.uint8, count=32 * n, offset=p).reshape(n, 32); p += 32 * n
    oe = O.OracleExtractor(nf, 1.2, 8, 20, 7)
    okps, odesc = oe(img)
    assert n == len(okps) and kps.tobytes() == okps.tobytes()
    assert np.array_equal(desc, odesc)
    L, = struct.unpack_from("<i", b, p); p += 4
    assert L == 8
    for l in range(L):
        lw, lh = struct.unpack_from("<ii", b, p); p += 8
        lev = np.frombuffer(b, dtype=np.uint8, count=lw * lh, offset=p).reshape(lh, lw); p += lw * lh
        assert (lw, lh) == oe.level_size(l)
        assert np.array_equal(lev, oe.level_padded(l)[19:-19, 19:-19]), "mvImagePyramid[%d]" % l
    nm, = struct.unpack_from("<i", b, p); p += 4
    m = np.frombuffer(b, dtype=np.int32, count=n, offset=p); p += 4 * n
    d01, = struct.unpack_from("<i", b, p); p += 4
    bi, bd, sd = O.match_bruteforce(odesc, odesc)
    acc = (bd <= 50) & (bd.astype(np.float32) < np.float32(0.9) * sd.astype(np.float32))
    assert nm == int(acc.sum())
    assert np.array_equal(m, np.where(acc, bi, -1))
    assert d01 == O.descriptor_distance(odesc[0], odesc[1])
    # map-side searches through the C++ wrappers (Fuse x2, SearchBySim3) against the oracle on the same re-projection
    bounds = (0.0, float(w), 0.0, float(h))
    sf, _, _, isg = oe.scale_tables()
    xy = np.stack([okps["x"] + np.float32(1.0), okps["y"]], 1).astype(np.float32)
    ur, xr = np.full(n, -1, np.float32), np.full(n, -1, np.float32)
    for th, chi2 in [(3.0, True), (4.0, False)]:
        nf, = struct.unpack_from("<i", b, p); p += 4
        got = np.frombuffer(b, dtype=np.int32, count=n, offset=p); p += 4 * n
        on, obi = O.fuse(okps, odesc, ur, bounds, sf, isg, xy, xr, okps["octave"], odesc, th, chi2)
        assert nf == on and np.array_equal(got, obi)
        assert nf > n // 2
    ns, = struct.unpack_from("<i", b, p); p += 4
    got = np.frombuffer(b, dtype=np.int32, count=n, offset=p); p += 4 * n
    on, om = O.search_by_sim3(okps, odesc, okps, odesc, bounds, sf, xy, okps["octave"], odesc, xy, okps["octave"], odesc, 7.5)
    assert ns == on and np.array_equal(got, om)
    assert ns > n // 2
    # Frame::ComputeBoW through include/ORBVocabulary.h (text loader + transform, levelsup 4)
    (obid, obval), (ofn, ofo, ofi) = O.OracleVocabulary(*voc).transform(odesc, 4)
    nb, = struct.unpack_from("<i", b, p); p += 4
    assert nb == len(obid) and nb > 100
    rec = np.frombuffer(b, dtype=np.dtype([("id", "<u4"), ("v", "<f8")]), count=nb, offset=p); p += 12 * nb
    assert np.array_equal(rec["id"], obid) and rec["v"].tobytes() == obval.tobytes()
    nfv, = struct.unpack_from("<i", b, p); p += 4
    assert nfv == len(ofn)
    for a in range(nfv):
        node, cnt = struct.unpack_from("<Ii", b, p); p += 8
        idx = np.frombuffer(b, dtype=np.uint32, count=cnt, offset=p); p += 4 * cnt
        assert node == ofn[a] and np.array_equal(idx, ofi[ofo[a]:ofo[a + 1]])
    assert p == len(b)


@pytest.mark.gpu
@pytest.mark.parametrize("opencv_shaped", [False, True], ids=["bundled_shim", "opencv_shaped_headers"])
def test_reference_signatures_equal_frameview_path(tmp_path, opencv_shaped):
    """SearchByProjection(Frame&, vpMapPoints, th), SearchByProjection(Current, Last, th, bMono), SearchByBoW x2 and
    SearchForInitialization(Frame&, Frame&, ...) on stand-in SLAM types give what the FrameView overloads (checked against
    the oracle in test_searches.py) give on the same inputs; the program exits non-zero on any difference."""
    exe = _build_native(tmp_path, "reference_api_test.cpp", "refapi", opencv_shaped)
    # two views of one scene: the second is the first moved by (3, 1) pixels with fresh noise, so every windowed search has work
    L = synth.rects(640, 480, 2100)
    rng = np.random.default_rng(2100)
    R = np.clip(np.roll(L, (1, 3), axis=(0, 1)).astype(np.int16) + rng.integers(-3, 4, L.shape), 0, 255).astype(np.uint8)
    (tmp_path / "a.raw").write_bytes(L.tobytes())
    (tmp_path / "b.raw").write_bytes(R.tobytes())
    r = subprocess.run([exe, str(tmp_path / "a.raw"), str(tmp_path / "b.raw"), "640", "480", "1000"], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.startswith("ok"), (r.returncode, r.stdout, r.stderr[-2000:])


@pytest.mark.gpu
def test_dropin_opencv_shaped_build_matches_oracle(tmp_path):
    """The ORBX_USE_OPENCV branch of the headers, run: same outputs as the bundled-shim build."""
    exe = _build_native(tmp_path, "dropin_test.cpp", "dropin_cv", True)
    img = synth.rects(640, 480, 1000)
    (tmp_path / "in.raw").write_bytes(img.tobytes())
    subprocess.check_call([exe, str(tmp_path / "in.raw"), "640", "480", "1000", str(tmp_path / "out.bin")])
    b = (tmp_path / "out.bin").read_bytes()
    n, = struct.unpack_from("<i", b, 0)
    okps, odesc = O.OracleExtractor(1000, 1.2, 8, 20, 7)(img)
    assert n == len(okps) and b[4:4 + 28 * n] == okps.tobytes() and b[4 + 28 * n:4 + 60 * n] == odesc.tobytes()
