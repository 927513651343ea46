"""SearchByProjection(Frame, MapPoints) (ORBmatcher.cc:45-129), SearchByProjection(CurrentFrame, LastFrame) (:1330-1472) and
both SearchByBoW (:159-288, :522-655).

CPU leg: the C++ oracle against an independent pure-Python restatement of the same reference loops (second opinion on the
orchestration: candidate order, occupancy updates, ratio / level rules, rotation histogram).  GPU leg: the C-ABI routines
(device candidate windows + Hamming distances, host replay) against the oracle, index for index."""
import numpy as np
import pytest

import oracle_py as O
import scenes

TH_HIGH, TH_LOW, HISTO = 100, 50, 30


def _dist(a, b):
    return int(np.unpackbits(a ^ b).sum())


def _three_maxima(hist):
    m1 = m2 = m3 = 0
    i1 = i2 = i3 = -1
    for i, h in enumerate(hist):
        s = len(h)
        if s > m1:
            m3, m2, m1, i3, i2, i1 = m2, m1, s, i2, i1, i
        elif s > m2:
            m3, m2, i3, i2 = m2, s, i2, i
        elif s > m3:
            m3, i3 = s, i
    if m2 < np.float32(0.1) * np.float32(m1):
        i2 = i3 = -1
    elif m3 < np.float32(0.1) * np.float32(m1):
        i3 = -1
    return i1, i2, i3


def _rot_bin(a1, a2):
    rot = np.float32(a1) - np.float32(a2)
    if rot < 0:
        rot = np.float32(rot + np.float32(360.0))
    b = int(np.floor(float(np.float32(rot * np.float32(1.0 / HISTO))) + 0.5))     # C round(): half away from zero, rot >= 0
    return 0 if b == HISTO else b


def py_search_by_projection(s, th, nnratio):
    k, d = s["kps"], s["desc"]
    occ = s["occupied"].copy()
    fm = np.full(len(k), -1, np.int32)
    n = 0
    for i in range(len(s["level"])):
        lv = int(s["level"][i])
        r = np.float32(2.5) if s["view_cos"][i] > 0.998 else np.float32(4.0)
        if th != 1.0:
            r = np.float32(r * np.float32(th))
        rad = np.float32(r * scenes.SF[lv])
        _, cand = O.grid_query_all(k, s["proj_xy"][i:i + 1], float(rad), lv - 1, lv, scenes.BOUNDS)
        best = best2 = 256
        bl = bl2 = bi = -1
        for j in cand:
            if occ[j]:
                continue
            if s["u_right"][j] > 0 and abs(np.float32(s["proj_xr"][i] - s["u_right"][j])) > rad:
                continue
            dd = _dist(s["mp_desc"][i], d[j])
            if dd < best:
                best2, best, bl2, bl, bi = best, dd, bl, int(k["octave"][j]), int(j)
            elif dd < best2:
                bl2, best2 = int(k["octave"][j]), dd
        if best <= TH_HIGH:
            if bl == bl2 and np.float32(best) > np.float32(nnratio) * np.float32(best2):
                continue
            fm[bi] = i
            occ[bi] = s["mp_has_obs"][i]
            n += 1
    return n, fm


def py_search_by_projection_frame(s, th, mode, check):
    k, d = s["kps"], s["desc"]
    occ = s["occupied"].copy()
    fm = np.full(len(k), -1, np.int32)
    hist = [[] for _ in range(HISTO)]
    n = 0
    for i in range(len(s["level"])):
        o = int(s["level"][i])
        rad = np.float32(np.float32(th) * scenes.SF[o])
        lo, hi = [(o - 1, o + 1), (o, -1), (0, o)][mode]
        _, cand = O.grid_query_all(k, s["proj_xy"][i:i + 1], float(rad), lo, hi, scenes.BOUNDS)
        best, bi = 256, -1
        for j in cand:
            if occ[j]:
                continue
            if s["u_right"][j] > 0 and abs(np.float32(s["proj_xr"][i] - s["u_right"][j])) > rad:
                continue
            dd = _dist(s["mp_desc"][i], d[j])
            if dd < best:
                best, bi = dd, int(j)
        if best <= TH_HIGH:
            fm[bi] = i
            occ[bi] = s["mp_has_obs"][i]
            n += 1
            if check:
                hist[_rot_bin(s["angle"][i], k["angle"][bi])].append(bi)
    if check:
        keep = _three_maxima(hist)
        for b in range(HISTO):
            if b not in keep:
                for j in hist[b]:
                    fm[j] = -1
                    n -= 1
    return n, fm


def py_search_by_bow(k1, d1, valid1, fv1, k2, d2, fv2, valid2, nnratio, check):
    variant = 0 if valid2 is None else 1
    out = np.full(len(k2) if variant == 0 else len(k1), -1, np.int32)
    taken = np.zeros(len(k2), bool)
    hist = [[] for _ in range(HISTO)]
    n = 0
    nodes2 = {int(nd): b for b, nd in enumerate(fv2[0])}
    for a, nd in enumerate(fv1[0]):
        if int(nd) not in nodes2:
            continue
        b = nodes2[int(nd)]
        for f1 in fv1[2][fv1[1][a]:fv1[1][a + 1]]:
            if not valid1[f1]:
                continue
            best = best2 = 256
            bi = -1
            for f2 in fv2[2][fv2[1][b]:fv2[1][b + 1]]:
                if taken[f2] or (variant == 1 and not valid2[f2]):
                    continue
                dd = _dist(d1[f1], d2[f2])
                if dd < best:
                    best2, best, bi = best, dd, int(f2)
                elif dd < best2:
                    best2 = dd
            low = best <= TH_LOW if variant == 0 else best < TH_LOW
            if low and np.float32(best) < np.float32(nnratio) * np.float32(best2):
                taken[bi] = True
                if variant == 0:
                    out[bi] = f1
                else:
                    out[f1] = bi
                if check:
                    hist[_rot_bin(k1["angle"][f1], k2["angle"][bi])].append(bi if variant == 0 else int(f1))
                n += 1
    if check:
        keep = _three_maxima(hist)
        for bb in range(HISTO):
            if bb not in keep:
                for j in hist[bb]:
                    out[j] = -1
                    n -= 1
    return n, out


# ------------------------------------------------------------------------------------------------ CPU: oracle vs Python
@pytest.fixture(scope="module")
def proj():
    return scenes.projection_scene(1000)


@pytest.fixture(scope="module")
def bow():
    return scenes.bow_scene(2000)


def _args(s):
    return (s["kps"], s["desc"], s["u_right"], s["occupied"], scenes.BOUNDS, scenes.SF, s["proj_xy"], s["proj_xr"], s["level"])


@pytest.mark.parametrize("th,nnratio", [(1.0, 0.8), (3.0, 0.8), (5.0, 0.6)])
def test_oracle_search_by_projection_vs_python(proj, th, nnratio):
    n, fm = O.search_by_projection(*_args(proj), proj["view_cos"], proj["mp_desc"], proj["mp_has_obs"], th, nnratio)
    pn, pfm = py_search_by_projection(proj, th, nnratio)
    assert n == pn and np.array_equal(fm, pfm)
    assert n > 50


@pytest.mark.parametrize("th,mode,check", [(15.0, 0, True), (7.0, 1, True), (15.0, 2, False)])
def test_oracle_search_by_projection_frame_vs_python(proj, th, mode, check):
    n, fm = O.search_by_projection_frame(*_args(proj), proj["angle"], proj["mp_desc"], proj["mp_has_obs"], th, mode, check)
    pn, pfm = py_search_by_projection_frame(proj, th, mode, check)
    assert n == pn and np.array_equal(fm, pfm)
    assert n > 50


@pytest.mark.parametrize("variant,nnratio,check", [(0, 0.7, True), (1, 0.8, True), (0, 0.9, False)])
def test_oracle_search_by_bow_vs_python(bow, variant, nnratio, check):
    v2 = None if variant == 0 else bow["valid2"]
    n, mt = O.search_by_bow(bow["k1"], bow["d1"], bow["valid1"], bow["fv1"], bow["k2"], bow["d2"], bow["fv2"], v2, nnratio, check)
    pn, pmt = py_search_by_bow(bow["k1"], bow["d1"], bow["valid1"], bow["fv1"], bow["k2"], bow["d2"], bow["fv2"], v2, nnratio, check)
    assert n == pn and np.array_equal(mt, pmt)
    assert n > 50


# ------------------------------------------------------------------------------------------------ GPU: C ABI vs oracle
@pytest.mark.gpu
@pytest.mark.parametrize("seed", [1000, 1001])
@pytest.mark.parametrize("th,nnratio", [(1.0, 0.8), (3.0, 0.8), (5.0, 0.6)])
def test_cuda_search_by_projection(seed, th, nnratio):
    import orbx
    s = scenes.projection_scene(seed)
    m = orbx.ORBmatcher(nnratio, True)
    n, fm = m.SearchByProjection(*_args(s), s["view_cos"], s["mp_desc"], s["mp_has_obs"], th)
    on, ofm = O.search_by_projection(*_args(s), s["view_cos"], s["mp_desc"], s["mp_has_obs"], th, nnratio)
    assert n == on and np.array_equal(fm, ofm)
    assert n > 50


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [1000, 1001])
@pytest.mark.parametrize("th,mode,check", [(15.0, 0, True), (7.0, 1, True), (15.0, 2, False), (30.0, 0, True)])
def test_cuda_search_by_projection_frame(seed, th, mode, check):
    import orbx
    s = scenes.projection_scene(seed)
    m = orbx.ORBmatcher(0.9, check)
    n, fm = m.SearchByProjectionFrame(*_args(s), s["angle"], s["mp_desc"], s["mp_has_obs"], th, mode)
    on, ofm = O.search_by_projection_frame(*_args(s), s["angle"], s["mp_desc"], s["mp_has_obs"], th, mode, check)
    assert n == on and np.array_equal(fm, ofm)
    assert n > 50


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [2000, 2001])
@pytest.mark.parametrize("variant,nnratio,check", [(0, 0.7, True), (1, 0.8, True), (0, 0.9, False), (1, 0.75, False)])
def test_cuda_search_by_bow(seed, variant, nnratio, check):
    import orbx
    b = scenes.bow_scene(seed)
    m = orbx.ORBmatcher(nnratio, check)
    for (k2, d2, fv2, v2) in [(b["k2"], b["d2"], b["fv2"], b["valid2"]), (b["kR"], b["dR"], b["fvR"], b["validR"])]:
        vv = None if variant == 0 else v2
        n, mt = m.SearchByBoW(b["k1"], b["d1"], b["valid1"], b["fv1"], k2, d2, fv2, vv)
        on, omt = O.search_by_bow(b["k1"], b["d1"], b["valid1"], b["fv1"], k2, d2, fv2, vv, nnratio, check)
        assert n == on and np.array_equal(mt, omt)
    assert n >= 0


@pytest.mark.gpu
@pytest.mark.parametrize("variant", [0, 1])
def test_cuda_search_by_bow_when_the_best_candidates_are_taken(variant):
    """The device returns the four smallest distances of every query; ten queries with one and the same descriptor compete
    for twelve near-copies of it, so from the fifth on all four are already matched and the bucket is rescanned on the host."""
    import orbx
    b = dict(scenes.bow_scene(2002))
    rng = np.random.default_rng(5)
    d1, d2 = b["d1"].copy(), b["d2"].copy()
    v1, v2 = b["valid1"].copy(), b["valid2"].copy()
    D = d1[17].copy()
    for i in range(40, 50):
        d1[i], v1[i] = D, 1
    for j in range(200, 212):
        d2[j] = D
        v2[j] = 1
        flip = rng.choice(np.arange(8, 256), size=j - 200, replace=False)      # 0..11 flipped bits, first byte (the node) kept
        for bit in flip:
            d2[j, bit >> 3] ^= np.uint8(1 << (bit & 7))
    b.update(d1=d1, d2=d2, valid1=v1, valid2=v2, fv1=scenes.feature_vector(d1), fv2=scenes.feature_vector(d2))
    for nnratio, check in [(0.99, False), (0.8, True)]:
        m = orbx.ORBmatcher(nnratio, check)
        vv = None if variant == 0 else b["valid2"]
        n, mt = m.SearchByBoW(b["k1"], b["d1"], b["valid1"], b["fv1"], b["k2"], b["d2"], b["fv2"], vv)
        on, omt = O.search_by_bow(b["k1"], b["d1"], b["valid1"], b["fv1"], b["k2"], b["d2"], b["fv2"], vv, nnratio, check)
        assert n == on and np.array_equal(mt, omt)
    assert n > 50


@pytest.mark.gpu
def test_cuda_searches_empty_inputs():
    import orbx
    s = scenes.projection_scene(1000)
    m = orbx.ORBmatcher(0.8, True)
    e = lambda dt, *shape: np.zeros(shape, dt)
    n, fm = m.SearchByProjection(s["kps"], s["desc"], s["u_right"], s["occupied"], scenes.BOUNDS, scenes.SF, e(np.float32, 0, 2), e(np.float32, 0),
                                 e(np.int32, 0), e(np.float32, 0), e(np.uint8, 0, 32), e(np.uint8, 0), 1.0)
    assert n == 0 and (fm == -1).all()
