"""The oracle's Frame-level restatements against the UNMODIFIED reference src/Frame.cc and the vendored DBoW2 (oracle/_ref/libframe_ref.so:
Frame.cc, ORBmatcher.cc, ORBextractor.cc, Thirdparty/DBoW2 compiled where they lie; oracle/ref_shim/frame_ref_api.cpp).

Every Frame here is built by the reference's own constructor and the numbers compared are the members it leaves behind:
  * Frame::ComputeStereoMatches (Frame.cc:481-655)          == oracle stereo_matches          
  * Frame::UndistortKeyPoints / ComputeImageBounds (:419-479) == oracle undistort_keypoints / image_bounds
  * Frame::ComputeStereoFromRGBD (:658-679)                   == oracle stereo_from_rgbd
  * Frame::AssignFeaturesToGrid / GetFeaturesInArea (:230-245, :342-395) == oracle grid
  * ORBVocabulary::transform via Frame::ComputeBoW (:410-417; TemplatedVocabulary.h:1140-1270, loadFromTextFile :1351-1436,
    BowVector.cpp, FeatureVector.cpp, ScoringObject.cpp) == oracle vocabulary
The GPU suite holds the CUDA path equal to the same oracle functions (tests/test_gpu_frame_matchers.py, test_frame_glue.py,
test_vocabulary.py, test_gpu_device_frame.py), which closes product == oracle == reference for these rows.  CPU only: no GPU needed."""
import os

import numpy as np
import pytest

import oracle_py as O
import ref_frame_py as RF
import scenes
from orbx import synth

pytestmark = pytest.mark.skipif(not RF.available(), reason="oracle/_ref/libframe_ref.so is not built and /root/reference is not here")


def _K(w, h, fx=520.0, fy=515.0):
    return np.array([fx, fy, w * 0.5 - 3.0, h * 0.5 + 2.0], np.float32)


# ---------------------------------------------------------------------------------------------- ComputeStereoMatches
@pytest.mark.parametrize("w,h,nf,fx,bf,seed", [(640, 480, 1000, 520.0, 40.0, 1000), (752, 480, 1200, 435.2, 47.9, 1001), (1241, 376, 2000, 718.9, 386.1, 1002),
                                               (640, 480, 1000, 520.0, 10.0, 1003), (320, 240, 500, 260.0, 20.0, 1004)])
def test_compute_stereo_matches_is_the_reference(w, h, nf, fx, bf, seed):
    left, right = synth.stereo_rects(w, h, seed)[:2]
    F = RF.RefFrame("stereo", left, _K(w, h, fx, fx), np.zeros(4), bf=bf, im_right=right, nfeatures=nf)
    assert F.N > nf // 2 and len(F.keys_right) > nf // 2
    assert F.mb == np.float32(np.float32(bf) / np.float32(fx))                                  # Frame.cc:225
    n, uR, dep = O.stereo_matches(F.keys, F.desc, F.keys_right, F.desc_right, F.pyramid(0), F.pyramid(1), F.sf, F.isf, bf,
                                  float(np.float32(bf) / F.mb))                                 # minZ = mb, maxD = mbf / minZ (:511-513)
    assert uR.tobytes() == F.u_right.tobytes() and dep.tobytes() == F.depth_out.tobytes()
    assert n == int((F.u_right >= 0).sum())
    assert n > 40                                                                               # not vacuous


def test_compute_stereo_matches_without_right_keypoints_or_matches():
    w, h = 640, 480
    left = synth.rects(w, h, 77)
    flat = np.full((h, w), 90, np.uint8)                                                        # no corner at all on the right
    F = RF.RefFrame("stereo", left, _K(w, h), np.zeros(4), bf=40.0, im_right=flat)
    assert len(F.keys_right) == 0 and (F.u_right == -1).all() and (F.depth_out == -1).all()
    n, uR, dep = O.stereo_matches(F.keys, F.desc, F.keys_right, F.desc_right, F.pyramid(0), F.pyramid(1), F.sf, F.isf, 40.0, float(np.float32(40.0) / F.mb))
    assert n == 0 and uR.tobytes() == F.u_right.tobytes() and dep.tobytes() == F.depth_out.tobytes()


def test_stereo_pipeline_of_the_cpu_baseline():
    """bench.py's stereo CPU legs: extractors kept, one reference Frame per pair.  Which of several equally sized quadtree nodes is split
    first depends on heap addresses, so the counts are bounded, not pinned."""
    w, h = 752, 480
    left, right = synth.stereo_rects(w, h, 1001)[:2]
    p = RF.RefStereoPipeline([435.2, 435.2, w * 0.5, h * 0.5], 47.9, nfeatures=1200)
    F = RF.RefFrame("stereo", left, [435.2, 435.2, w * 0.5, h * 0.5], np.zeros(4), bf=47.9, im_right=right, nfeatures=1200)
    for _ in range(2):
        n, matched = p.run(left, right)
        assert abs(n - F.N) <= 40 and abs(matched - int((F.u_right >= 0).sum())) <= 60 and matched > 100


# ---------------------------------------------------------------------------------------------- undistortion, bounds, RGB-D
CAMERAS = [("tum1", (517.3, 516.5, 318.6, 255.3), (0.2624, -0.9531, -0.0054, 0.0026, 1.1633)),
           ("tum2", (520.9, 521.0, 325.1, 249.7), (0.2312, -0.7849, -0.0033, -0.0001, 0.9172)),
           ("four", (458.6, 457.3, 367.2, 248.4), (-0.2834, 0.0739, 0.00019, 1.76e-05)),
           ("none", (535.4, 539.2, 320.1, 247.6), (0.0, 0.0, 0.0, 0.0, 0.0))]


@pytest.mark.parametrize("name,K4,dist", CAMERAS)
def test_rgbd_frame_glue_is_the_reference(name, K4, dist):
    w, h, bf = 640, 480, 40.0
    im = synth.rects(w, h, 1200 + len(name))
    rng = np.random.default_rng(5)
    depth = rng.uniform(0.4, 9.0, (h, w)).astype(np.float32)
    depth[rng.random((h, w)) < 0.2] = 0.0                                                       # holes: no stereo coordinate there
    depth[rng.random((h, w)) < 0.02] = -1.0
    F = RF.RefFrame("rgbd", im, K4, dist, bf=bf, depth=depth)
    assert F.N > 500
    K4f = [np.float32(v) for v in K4]
    un = O.undistort_keypoints(F.keys, K4f, dist)
    assert un.tobytes() == F.keys_un.tobytes()
    assert (dist[0] == 0) == (F.keys_un.tobytes() == F.keys.tobytes())
    assert O.image_bounds(w, h, K4f, dist).tobytes() == F.bounds.tobytes()
    uR, dep = O.stereo_from_rgbd(F.keys, F.keys_un, depth, bf)
    assert uR.tobytes() == F.u_right.tobytes() and dep.tobytes() == F.depth_out.tobytes()
    assert 0 < int((dep > 0).sum()) < F.N
    # grid scale (Frame.cc:155-156)
    assert F.grid_inv[0] == np.float32(64) / (F.bounds[1] - F.bounds[0]) and F.grid_inv[1] == np.float32(48) / (F.bounds[3] - F.bounds[2])


# ---------------------------------------------------------------------------------------------- the feature grid
@pytest.mark.parametrize("name,K4,dist", [CAMERAS[0], CAMERAS[3]])
def test_grid_and_features_in_area_are_the_reference(name, K4, dist):
    w, h = 640, 480
    F = RF.RefFrame("mono", synth.rects(w, h, 1300), K4, dist, bf=40.0, nfeatures=1500)
    G = O.lib().orbo_grid_build(O._p(F.keys_un), F.N, *[float(b) for b in F.bounds])
    try:
        buf = np.zeros(F.N, np.int32)
        rng = np.random.default_rng(9)
        nonempty = 0
        for q in range(600):
            if q % 3 == 0:                              # around a keypoint, as the searches do
                i = int(rng.integers(0, F.N))
                x, y = float(F.keys_un["x"][i] + rng.uniform(-4, 4)), float(F.keys_un["y"][i] + rng.uniform(-4, 4))
            else:                                       # anywhere, including outside the image
                x, y = float(rng.uniform(-60, w + 60)), float(rng.uniform(-60, h + 60))
            r = float(rng.choice([1.0, 7.5, 15.0, 40.0, 100.0, 300.0]))
            lo, hi = [(-1, -1), (0, -1), (2, -1), (0, 3), (2, 5), (1, 7), (-1, 4)][int(rng.integers(0, 7))]
            n = O.lib().orbo_grid_query(G, O._p(F.keys_un), x, y, r, lo, hi, O._p(buf), len(buf))
            ref = F.features_in_area(x, y, r, lo, hi)
            assert buf[:n].tolist() == ref.tolist(), (x, y, r, lo, hi)          # same features in the same order
            nonempty += n > 0
        assert nonempty > 250
    finally:
        O.lib().orbo_grid_destroy(G)
    # AssignFeaturesToGrid: every feature sits in the cell PosInGrid gives (Frame.cc:230-245, :397-407), in index order
    total = 0
    for col in range(64):
        for row in range(48):
            cell = F.grid_cell(col, row)
            total += len(cell)
            assert cell.tolist() == sorted(cell.tolist())
            for i in cell[:2]:
                ok, px, py = F.pos_in_grid(float(F.keys_un["x"][i]), float(F.keys_un["y"][i]))
                assert ok and (px, py) == (col, row)
    outside = sum(not F.pos_in_grid(float(k["x"]), float(k["y"]))[0] for k in F.keys_un)
    assert total + outside == F.N


# ---------------------------------------------------------------------------------------------- the vocabulary
def _write_voc(path, voc, scoring=0, weighting=0, trailing_newline=False):
    k, L, parent, desc, weight = voc
    scenes.write_vocabulary_text(path, k, L, parent, desc, weight, scoring, weighting)
    if not trailing_newline:                            # see test_vocabulary_text_file_with_a_trailing_newline
        with open(path, "rb+") as f:
            f.seek(-1, os.SEEK_END)
            if f.read(1) == b"\n":
                f.seek(-1, os.SEEK_END)
                f.truncate()


def _depth_of(parent):
    d = np.zeros(len(parent), np.int32)
    for i in range(1, len(parent)):
        d[i] = d[parent[i]] + 1
    return d


@pytest.mark.parametrize("scoring,weighting", [(0, 0), (1, 0), (2, 0), (3, 0), (4, 0), (5, 0), (0, 1), (0, 2), (0, 3), (5, 3), (3, 1), (4, 2)])
def test_vocabulary_transform_is_dbow2(tmp_path, scoring, weighting):
    (k1, d1), (k2, d2) = scenes.frames(1000)
    feats = np.concatenate([d1, d2, np.random.default_rng(3).integers(0, 256, (300, 32)).astype(np.uint8)])
    voc = scenes.synthetic_vocabulary(3, 10, 4, shallow_leaf_prob=0.0, desc_pool=d1)            # complete tree: nid defined for every levelsup
    path = str(tmp_path / "voc.txt")
    _write_voc(path, voc, scoring, weighting)
    rv = RF.RefVocabulary(path)
    ov = O.OracleVocabulary(*voc, scoring, weighting)
    assert rv.size() == int((np.bincount(voc[2][1:], minlength=len(voc[2])) == 0).sum())
    for levelsup in (0, 1, 2, 4, 9):
        (b, v), fv = rv.transform(feats, levelsup)
        (ob, ovv), ofv = ov.transform(feats, levelsup)
        assert np.array_equal(b, ob) and v.tobytes() == ovv.tobytes()                           # BowVector: ids and the doubles, bit for bit
        for x, y in zip(fv, ofv):
            assert np.array_equal(x, y)                                                         # FeatureVector: nodes, offsets, feature indices
        assert len(b) > 200
    (b, v), fv = rv.transform(feats[:0], 4)
    assert len(b) == 0 and len(fv[0]) == 0


def test_vocabulary_with_shallow_leaves_where_the_reference_is_defined(tmp_path):
    """A leaf above level L - levelsup leaves *nid unset in the reference (TemplatedVocabulary.h:1166, :1259: `NodeId nid;` is read
    uninitialised); everything else must agree: the BowVector always, the FeatureVector for every feature whose word is deep enough."""
    (k1, d1), _ = scenes.frames(1001)
    voc = scenes.synthetic_vocabulary(5, 10, 4, shallow_leaf_prob=0.08, desc_pool=d1)
    k, L, parent, desc, weight = voc
    path = str(tmp_path / "voc.txt")
    _write_voc(path, voc)
    rv, ov = RF.RefVocabulary(path), O.OracleVocabulary(*voc)
    depth = _depth_of(parent)
    is_leaf = np.bincount(parent[1:], minlength=len(parent)) == 0
    leaves = np.flatnonzero(is_leaf)                                                            # word id = rank among the leaves, in node order
    for levelsup in (0, 1, 2, 3, 4):
        (b, v), (fn, fo, fi) = rv.transform(d1, levelsup)
        (ob, ovv), (ofn, ofo, ofi) = ov.transform(d1, levelsup)
        assert np.array_equal(b, ob) and v.tobytes() == ovv.tobytes()
        wid, w, nid = ov.transform_raw(d1, levelsup)
        defined = depth[leaves[wid]] >= L - levelsup
        node_ref = np.full(len(d1), -1, np.int64)
        for a in range(len(fn)):
            node_ref[fi[fo[a]:fo[a + 1]]] = fn[a]
        kept = w > 0
        assert ((node_ref >= 0) == kept).all()                                                  # stopped words are left out by both
        sel = kept & defined
        assert np.array_equal(node_ref[sel], nid[sel]) and sel.sum() > 300
        if levelsup >= L - 1:
            assert defined.all()
    assert (~(depth[leaves] >= L)).any()                                                        # the vocabulary does have shallow leaves


def test_frame_compute_bow_is_the_reference(tmp_path):
    (k1, d1), _ = scenes.frames(1002)
    voc = scenes.synthetic_vocabulary(7, 10, 5, shallow_leaf_prob=0.0, desc_pool=d1)            # L = 5, levelsup = 4 as Frame.cc:415 passes
    path = str(tmp_path / "voc.txt")
    _write_voc(path, voc)
    rv = RF.RefVocabulary(path)
    w, h = 640, 480
    F = RF.RefFrame("mono", synth.rects(w, h, 1400), _K(w, h), np.zeros(4), bf=40.0, voc=rv)
    (b, v), fv = F.compute_bow()
    (ob, ovv), ofv = O.OracleVocabulary(*voc).transform(F.desc, 4)
    assert np.array_equal(b, ob) and v.tobytes() == ovv.tobytes()
    for x, y in zip(fv, ofv):
        assert np.array_equal(x, y)
    assert len(b) > 300 and abs(v.sum() - 1.0) < 1e-9                                           # L1-normalised TF-IDF
    # DBoW2's L1 score of the vector with itself is 1 (ScoringObject.cpp), and below 1 against another frame
    F2 = RF.RefFrame("mono", synth.rects(w, h, 1401), _K(w, h), np.zeros(4), bf=40.0, voc=rv)
    b2 = F2.compute_bow()[0]
    assert abs(rv.score((b, v), (b, v)) - 1.0) < 1e-12 and 0.0 <= rv.score((b, v), b2) < 0.9


def test_vocabulary_text_file_with_a_trailing_newline(tmp_path):
    """saveToTextFile ends every line with endl, and loadFromTextFile loops `while(!f.eof())` (TemplatedVocabulary.h:1394): the empty
    last line becomes one more node.  Its `pid` is the first failed extraction (0: a child of the root); `nIsLeaf` and the weight are
    then never written (the stream is already in the fail state), so whether the node also becomes a word depends on an uninitialised
    int (:1411-1412), and FORB::fromString never writes its descriptor bytes (FORB.cpp:121-135).  If it is a word, its weight is the
    Node() default 0, i.e. a stopped word that contributes to neither vector.  The product's loader skips the empty line
    (include/ORBVocabulary.h); the comparisons above use files without the trailing newline, where the reference is defined."""
    (k1, d1), _ = scenes.frames(1000)
    voc = scenes.synthetic_vocabulary(3, 10, 3, shallow_leaf_prob=0.0, desc_pool=d1)
    a, b = str(tmp_path / "a.txt"), str(tmp_path / "b.txt")
    _write_voc(a, voc, trailing_newline=False)
    _write_voc(b, voc, trailing_newline=True)
    assert RF.RefVocabulary(a).size() == 1000
    assert RF.RefVocabulary(b).size() in (1000, 1001)


# ---------------------------------------------------------------------------------------------- the swap itself
# oracle/_ref/libframe_dropin*.so: the reference's UNMODIFIED Frame.cc / KeyFrame.cc / MapPoint.cc compiled against THIS REPOSITORY'S
# include/ORBextractor.h and include/ORBmatcher.h in place of the reference's headers (oracle/ref_shim/dropin/prelude_dropin.h) -- no
# ORBextractor.cc, no ORBmatcher.cc of the reference inside.  The reference's own Frame constructor then drives the product's extractor
# (operator(), GetLevels() .. GetInverseScaleSigmaSquares()), and the reference's own ComputeStereoMatches reads the product's mvImagePyramid.
STEREO_SHAPES = [(640, 480, 1000, 520.0, 40.0, 1000), (752, 480, 1200, 435.2, 47.9, 1001), (1241, 376, 2000, 718.9, 386.1, 1002)]


def _check_dropin_stereo_frame(libpath, w, h, nf, fx, bf, seed, keys_of):
    left, right = synth.stereo_rects(w, h, seed)[:2]
    F = RF.RefFrame("stereo", left, _K(w, h, fx, fx), np.zeros(4), bf=bf, im_right=right, nfeatures=nf, libpath=libpath)
    (kl, dl), (kr, dr) = keys_of(left), keys_of(right)
    assert F.keys.tobytes() == kl.tobytes() and F.desc.tobytes() == dl.tobytes()                # Frame::ExtractORB(0, .) ran the product's operator()
    assert F.keys_right.tobytes() == kr.tobytes() and F.desc_right.tobytes() == dr.tobytes()
    assert F.keys_un.tobytes() == F.keys.tobytes() and F.bounds.tolist() == [0.0, float(w), 0.0, float(h)]
    oe = O.OracleExtractor(nf, 1.2, 8, 20, 7)
    sf = oe.scale_tables()
    assert F.sf.tobytes() == sf[0].tobytes() and F.isf.tobytes() == sf[1].tobytes() and F.sigma2.tobytes() == sf[2].tobytes() and F.isigma2.tobytes() == sf[3].tobytes()
    # the reference's ComputeStereoMatches over the product's mvImagePyramid == the oracle over the same views
    n, uR, dep = O.stereo_matches(F.keys, F.desc, F.keys_right, F.desc_right, F.pyramid(0), F.pyramid(1), F.sf, F.isf, bf, float(np.float32(bf) / F.mb))
    assert uR.tobytes() == F.u_right.tobytes() and dep.tobytes() == F.depth_out.tobytes() and n > 40
    # ... and the product's mvImagePyramid views are the oracle's pyramid levels
    oe(left)
    for l, lv in enumerate(F.pyramid(0)):
        assert np.array_equal(lv, oe.level_padded(l)[19:-19, 19:-19]), l
    # the grid the reference built over the product's keypoints
    total = sum(len(F.grid_cell(c, r)) for c in range(64) for r in range(48))
    assert total == F.N - sum(not F.pos_in_grid(float(k["x"]), float(k["y"]))[0] for k in F.keys_un)
    return F, n, uR, dep


@pytest.mark.parametrize("w,h,nf,fx,bf,seed", STEREO_SHAPES)
def test_reference_frame_built_on_the_drop_in_extractor_cpu(w, h, nf, fx, bf, seed):
    if not RF.dropin_available(cpu=True):
        pytest.skip("oracle/_ref/libframe_dropin_cpu.so is not built and the reference tree is not here")
    oe = O.OracleExtractor(nf, 1.2, 8, 20, 7)
    _check_dropin_stereo_frame(RF.DROPIN_CPU_PATH, w, h, nf, fx, bf, seed, lambda im: oe(im))


def test_reference_rgbd_frame_built_on_the_drop_in_extractor_cpu():
    if not RF.dropin_available(cpu=True):
        pytest.skip("oracle/_ref/libframe_dropin_cpu.so is not built and the reference tree is not here")
    name, K4, dist = CAMERAS[0]
    w, h, bf = 640, 480, 40.0
    im = synth.rects(w, h, 1500)
    depth = np.random.default_rng(6).uniform(0.4, 9.0, (h, w)).astype(np.float32)
    F = RF.RefFrame("rgbd", im, K4, dist, bf=bf, depth=depth, libpath=RF.DROPIN_CPU_PATH)
    k, d = O.OracleExtractor(1000, 1.2, 8, 20, 7)(im)
    assert F.keys.tobytes() == k.tobytes() and F.desc.tobytes() == d.tobytes()
    K4f = [np.float32(v) for v in K4]
    assert O.undistort_keypoints(F.keys, K4f, dist).tobytes() == F.keys_un.tobytes()
    uR, dep = O.stereo_from_rgbd(F.keys, F.keys_un, depth, bf)
    assert uR.tobytes() == F.u_right.tobytes() and dep.tobytes() == F.depth_out.tobytes()


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,nf,fx,bf,seed", STEREO_SHAPES)
def test_reference_frame_built_on_the_cuda_extractor(w, h, nf, fx, bf, seed):
    """The reference's Frame constructor and ComputeStereoMatches (its own CPU code) around the CUDA extractor, through the drop-in header;
    the stereo association it computes equals the product's own CUDA ComputeStereoMatches on the same frame."""
    if not RF.dropin_available(cpu=False):
        pytest.skip("oracle/_ref/libframe_dropin.so was not built (make -C oracle _ref_tests needs the reference tree)")
    import orbx
    exL, exR = orbx.ORBextractor(nf, 1.2, 8, 20, 7), orbx.ORBextractor(nf, 1.2, 8, 20, 7)
    left, right = synth.stereo_rects(w, h, seed)[:2]
    F, n, uR, dep = _check_dropin_stereo_frame(RF.DROPIN_PATH, w, h, nf, fx, bf, seed, lambda im: orbx.ORBextractor(nf, 1.2, 8, 20, 7)(im))
    kl, dl = exL(left)
    kr, dr = exR(right)
    m = orbx.ORBmatcher()
    n2, uR2, dep2 = m.ComputeStereoMatches(exL, exR, kl, dl, kr, dr, bf, float(np.float32(bf) / F.mb))
    assert n2 == n and uR2.tobytes() == uR.tobytes() and dep2.tobytes() == dep.tobytes()


def _check_dropin_bow(libpath, tmp_path):
    """Frame::ComputeBoW of the reference (Frame.cc:410-417) calling the drop-in ORBVocabulary::transform with its own DBoW2 members."""
    (k1, d1), _ = scenes.frames(1002)
    voc = scenes.synthetic_vocabulary(7, 10, 5, shallow_leaf_prob=0.0, desc_pool=d1)
    path = str(tmp_path / "voc.txt")
    _write_voc(path, voc, trailing_newline=True)                                                # as saveToTextFile leaves it
    pv = RF.RefVocabulary(path, libpath=libpath)                                                # the drop-in loadFromTextFile
    assert pv.size() == int((np.bincount(voc[2][1:], minlength=len(voc[2])) == 0).sum())
    w, h = 640, 480
    F = RF.RefFrame("mono", synth.rects(w, h, 1400), _K(w, h), np.zeros(4), bf=40.0, voc=pv, libpath=libpath)
    (b, v), fv = F.compute_bow()
    (ob, ovv), ofv = O.OracleVocabulary(*voc).transform(F.desc, 4)
    assert np.array_equal(b, ob) and v.tobytes() == ovv.tobytes() and len(b) > 300
    for x, y in zip(fv, ofv):
        assert np.array_equal(x, y)
    return F, (b, v), fv, path


def test_reference_frame_compute_bow_on_the_drop_in_vocabulary_cpu(tmp_path):
    if not RF.dropin_available(cpu=True):
        pytest.skip("oracle/_ref/libframe_dropin_cpu.so is not built and the reference tree is not here")
    F, bow, fv, path = _check_dropin_bow(RF.DROPIN_CPU_PATH, tmp_path)
    # the same frame through the reference's own DBoW2 (libframe_ref.so): identical vectors
    _write_voc(path, scenes.synthetic_vocabulary(7, 10, 5, shallow_leaf_prob=0.0, desc_pool=scenes.frames(1002)[0][1]))
    F2 = RF.RefFrame("mono", synth.rects(640, 480, 1400), _K(640, 480), np.zeros(4), bf=40.0, voc=RF.RefVocabulary(path))
    if F2.desc.tobytes() == F.desc.tobytes():                   # the reference extractor's tie-breaks follow the heap; usually the same set here
        (b2, v2), fv2 = F2.compute_bow()
        assert np.array_equal(bow[0], b2) and bow[1].tobytes() == v2.tobytes()
        for x, y in zip(fv, fv2):
            assert np.array_equal(x, y)


@pytest.mark.gpu
def test_reference_frame_compute_bow_on_the_cuda_vocabulary(tmp_path):
    if not RF.dropin_available(cpu=False):
        pytest.skip("oracle/_ref/libframe_dropin.so was not built (make -C oracle _ref_tests needs the reference tree)")
    _check_dropin_bow(RF.DROPIN_PATH, tmp_path)


@pytest.mark.parametrize("scoring", [0, 1, 2, 3, 4, 5])
def test_drop_in_vocabulary_score_and_binary_file_are_dbow2(tmp_path, scoring):
    """score(v1, v2) (ScoringObject.cpp) and loadFromBinaryFile (TemplatedVocabulary.h:1467-1511) of include/ORBVocabulary.h against the
    vendored DBoW2: the binary file is written by DBoW2's own saveToBinaryFile."""
    if not RF.dropin_available(cpu=True):
        pytest.skip("oracle/_ref/libframe_dropin_cpu.so is not built and the reference tree is not here")
    (k1, d1), (k2, d2) = scenes.frames(1003)
    voc = scenes.synthetic_vocabulary(11, 10, 4, shallow_leaf_prob=0.0, desc_pool=d1)
    txt, binf = str(tmp_path / "voc.txt"), str(tmp_path / "voc.bin")
    _write_voc(txt, voc, scoring, 0)
    dbow = RF.RefVocabulary(txt)
    dbow.save_binary(binf)
    mine_txt = RF.RefVocabulary(txt, libpath=RF.DROPIN_CPU_PATH)
    mine_bin = RF.RefVocabulary(binf, libpath=RF.DROPIN_CPU_PATH, binary=True)
    dbow_bin = RF.RefVocabulary(binf, binary=True)
    nwords = int((np.bincount(voc[2][1:], minlength=len(voc[2])) == 0).sum())
    assert mine_txt.size() == mine_bin.size() == dbow.size() == nwords
    assert dbow_bin.size() in (nwords, nwords + 1)           # DBoW2's reader files the last record twice (see include/ORBVocabulary.h)
    # the binary format keeps the weights as float
    k, L, parent, desc, weight = voc
    ov32 = O.OracleVocabulary(k, L, parent, desc, weight.astype(np.float32).astype(np.float64), scoring, 0)
    for feats in (d1, d2):
        (b, v), fv = mine_bin.transform(feats, 4)
        (rb, rv), rfv = dbow_bin.transform(feats, 4)
        (ob, ovv), ofv = ov32.transform(feats, 4)
        assert np.array_equal(b, ob) and v.tobytes() == ovv.tobytes() and np.array_equal(b, rb) and v.tobytes() == rv.tobytes()
        for x, y, z in zip(fv, ofv, rfv):
            assert np.array_equal(x, y) and np.array_equal(x, z)
    # scores: the two frames against each other and themselves, both ways, and against an empty vector
    va, vb = dbow.transform(d1, 4)[0], dbow.transform(d2, 4)[0]
    empty = (np.zeros(0, np.int32), np.zeros(0, np.float64))
    sub = (va[0][::3], va[1][::3])
    seen = set()
    for x, y in [(va, vb), (vb, va), (va, va), (va, sub), (sub, vb), (va, empty), (empty, vb)]:
        r, m = dbow.score(x, y), mine_txt.score(x, y)
        assert np.float64(r).tobytes() == np.float64(m).tobytes(), (scoring, r, m)
        seen.add(round(r, 6))
    assert len(seen) >= 3
