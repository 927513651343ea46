"""The LocalMapping / LoopClosing / relocalisation searches: Fuse x2 (ORBmatcher.cc:827-1102), SearchBySim3 (:1104-1328),
SearchByProjection(KF, Scw) (:290-403), SearchByProjection(Frame, KF) (:1474-1601), SearchForTriangulation (:657-825).

CPU leg: the C++ oracle against an independent pure-Python restatement of the same reference loops.  GPU leg: the C-ABI
routines (one-kernel window + gate + argmin for Fuse / Sim3; device distances + host replay for the greedy three) against the
oracle, index for index."""
import numpy as np
import pytest

import oracle_py as O
import scenes
from test_searches import _dist, _rot_bin, _three_maxima

TH_HIGH, TH_LOW, HISTO = 100, 50, 30
f32 = np.float32


def py_area_best(k, d, u, v, radius, level, dmp, u_right=None, ur=None, inv_sigma2=None, skip=None):
    """KeyFrame::GetFeaturesInArea + the candidate loop of Fuse / SearchBySim3."""
    _, cand = O.grid_query_all(k, np.array([[u, v]], f32), float(radius), -1, -1, scenes.BOUNDS)
    best, bi = 256, -1
    for j in cand:
        if skip is not None and skip[j]:
            continue
        lv = int(k["octave"][j])
        if lv < level - 1 or lv > level:
            continue
        if inv_sigma2 is not None:
            ex, ey = f32(u) - k["x"][j], f32(v) - k["y"][j]
            if u_right[j] >= 0:
                er = f32(ur) - u_right[j]
                e2 = f32(f32(f32(ex * ex) + f32(ey * ey)) + f32(er * er))
                if float(f32(e2 * inv_sigma2[lv])) > 7.8:
                    continue
            else:
                e2 = f32(f32(ex * ex) + f32(ey * ey))
                if float(f32(e2 * inv_sigma2[lv])) > 5.99:
                    continue
        dd = _dist(dmp, d[j])
        if dd < best:
            best, bi = dd, int(j)
    return bi, best


def py_fuse(s, th, chi2):
    out = np.full(len(s["level"]), -1, np.int32)
    n = 0
    for i in range(len(s["level"])):
        lv = int(s["level"][i])
        bi, best = py_area_best(s["kps"], s["desc"], s["proj_xy"][i, 0], s["proj_xy"][i, 1], f32(f32(th) * scenes.SF[lv]), lv, s["mp_desc"][i],
                                s["u_right"], s["proj_ur"][i], scenes.INV_SIGMA2 if chi2 else None)
        if best <= TH_LOW:
            out[i] = bi
            n += 1
    return n, out


def py_search_by_sim3(s, th):
    def one_way(kq_n, pxy, lvl, mpd, kt, dt):
        m = np.full(kq_n, -1, np.int32)
        for i in range(kq_n):
            lv = int(lvl[i])
            if lv < 0:
                continue
            bi, best = py_area_best(kt, dt, pxy[i, 0], pxy[i, 1], f32(f32(th) * scenes.SF[lv]), lv, mpd[i])
            if best <= TH_HIGH:
                m[i] = bi
        return m
    m1 = one_way(len(s["k1"]), s["proj12_xy"], s["level12"], s["mp1_desc"], s["k2"], s["d2"])
    m2 = one_way(len(s["k2"]), s["proj21_xy"], s["level21"], s["mp2_desc"], s["k1"], s["d1"])
    out = np.full(len(s["k1"]), -1, np.int32)
    n = 0
    for i1 in range(len(s["k1"])):
        if m1[i1] >= 0 and m2[m1[i1]] == i1:
            out[i1] = m1[i1]
            n += 1
    return n, out


def py_search_by_projection_sim3(s, th):
    k, d = s["kps"], s["desc"]
    matched = s["occupied"].copy()
    fm = np.full(len(k), -1, np.int32)
    n = 0
    for i in range(len(s["level"])):
        lv = int(s["level"][i])
        bi, best = py_area_best(k, d, s["proj_xy"][i, 0], s["proj_xy"][i, 1], f32(f32(th) * scenes.SF[lv]), lv, s["mp_desc"][i], skip=matched)
        if best <= TH_LOW:
            fm[bi] = i
            matched[bi] = 1
            n += 1
    return n, fm


def py_search_by_projection_kf(s, th, orb_dist, check):
    k, d = s["kps"], s["desc"]
    occ = s["occupied"].copy()
    fm = np.full(len(k), -1, np.int32)
    hist = [[] for _ in range(HISTO)]
    n = 0
    for i in range(len(s["level"])):
        lv = int(s["level"][i])
        _, cand = O.grid_query_all(k, s["proj_xy"][i:i + 1], float(f32(f32(th) * scenes.SF[lv])), lv - 1, lv + 1, scenes.BOUNDS)
        best, bi = 256, -1
        for j in cand:
            if occ[j]:
                continue
            dd = _dist(s["mp_desc"][i], d[j])
            if dd < best:
                best, bi = dd, int(j)
        if best <= orb_dist:
            fm[bi] = i
            occ[bi] = 1
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


def py_search_for_triangulation(s, only_stereo, check):
    k1, d1, k2, d2 = s["k1"], s["d1"], s["k2"], s["d2"]
    F = s["F12"].astype(f32)
    ex, ey = s["epipole"]
    out = np.full(len(k1), -1, np.int32)
    matched2 = np.zeros(len(k2), bool)
    hist = [[] for _ in range(HISTO)]
    n = 0
    nodes2 = {int(nd): b for b, nd in enumerate(s["fv2"][0])}
    for a, nd in enumerate(s["fv1"][0]):
        if int(nd) not in nodes2:
            continue
        b = nodes2[int(nd)]
        for i1 in s["fv1"][2][s["fv1"][1][a]:s["fv1"][1][a + 1]]:
            if s["has_mp1"][i1]:
                continue
            st1 = s["u_right1"][i1] >= 0
            if only_stereo and not st1:
                continue
            x1, y1 = k1["x"][i1], k1["y"][i1]
            best, bi = TH_LOW, -1
            for i2 in s["fv2"][2][s["fv2"][1][b]:s["fv2"][1][b + 1]]:
                if matched2[i2] or s["has_mp2"][i2]:
                    continue
                st2 = s["u_right2"][i2] >= 0
                if only_stereo and not st2:
                    continue
                dd = _dist(d1[i1], d2[i2])
                if dd > TH_LOW or dd > best:
                    continue
                x2, y2, o2 = k2["x"][i2], k2["y"][i2], int(k2["octave"][i2])
                if not st1 and not st2:
                    dx, dy = f32(ex) - x2, f32(ey) - y2
                    if f32(f32(dx * dx) + f32(dy * dy)) < f32(f32(100) * scenes.SF[o2]):
                        continue
                la = f32(f32(f32(x1 * F[0, 0]) + f32(y1 * F[1, 0])) + F[2, 0])
                lb = f32(f32(f32(x1 * F[0, 1]) + f32(y1 * F[1, 1])) + F[2, 1])
                lc = f32(f32(f32(x1 * F[0, 2]) + f32(y1 * F[1, 2])) + F[2, 2])
                num = f32(f32(f32(la * x2) + f32(lb * y2)) + lc)
                den = f32(f32(la * la) + f32(lb * lb))
                if den == 0:
                    continue
                dsqr = f32(f32(num * num) / den)
                if float(dsqr) < 3.84 * float(scenes.SIGMA2[o2]):
                    bi, best = int(i2), dd
            if bi >= 0:
                out[i1] = bi
                matched2[bi] = True
                n += 1
                if check:
                    hist[_rot_bin(k1["angle"][i1], k2["angle"][bi])].append(int(i1))
    if check:
        keep = _three_maxima(hist)
        for bb in range(HISTO):
            if bb not in keep:
                for j in hist[bb]:
                    matched2[out[j]] = False
                    out[j] = -1
                    n -= 1
    return n, out


# ------------------------------------------------------------------------------------------------ CPU: oracle vs Python
@pytest.fixture(scope="module")
def fuse_s():
    return scenes.fuse_scene(1000)


@pytest.fixture(scope="module")
def proj_s():
    return scenes.projection_scene(1000)


def _fuse_args(s):
    return (s["kps"], s["desc"], s["u_right"], scenes.BOUNDS, scenes.SF, scenes.INV_SIGMA2, s["proj_xy"], s["proj_ur"], s["level"], s["mp_desc"])


def _tri_args(s):
    return (s["k1"], s["d1"], s["has_mp1"], s["u_right1"], s["fv1"], s["k2"], s["d2"], s["has_mp2"], s["u_right2"], s["fv2"], s["F12"], s["epipole"],
            scenes.SF, scenes.SIGMA2)


def _sim3_args(s):
    return (s["k1"], s["d1"], s["k2"], s["d2"], scenes.BOUNDS, scenes.SF, s["proj12_xy"], s["level12"], s["mp1_desc"], s["proj21_xy"], s["level21"],
            s["mp2_desc"])


@pytest.mark.parametrize("th,chi2", [(3.0, True), (4.0, False), (2.5, True)])
def test_oracle_fuse_vs_python(fuse_s, th, chi2):
    n, bi = O.fuse(*_fuse_args(fuse_s), th, chi2)
    pn, pbi = py_fuse(fuse_s, th, chi2)
    assert n == pn and np.array_equal(bi, pbi)
    assert n > 50
    if chi2:        # the gate must actually decide something in this scene
        n0, _ = O.fuse(*_fuse_args(fuse_s), th, False)
        assert n0 > n


def test_oracle_search_by_sim3_vs_python():
    s = scenes.sim3_scene(1000)
    n, mt = O.search_by_sim3(*_sim3_args(s), 7.5)
    pn, pmt = py_search_by_sim3(s, 7.5)
    assert n == pn and np.array_equal(mt, pmt)
    assert n > 50


@pytest.mark.parametrize("th", [10.0, 4.0])
def test_oracle_search_by_projection_sim3_vs_python(proj_s, th):
    s = proj_s
    n, fm = O.search_by_projection_sim3(s["kps"], s["desc"], s["occupied"], scenes.BOUNDS, scenes.SF, s["proj_xy"], s["level"], s["mp_desc"], th)
    pn, pfm = py_search_by_projection_sim3(s, th)
    assert n == pn and np.array_equal(fm, pfm)
    assert n > 50


@pytest.mark.parametrize("th,orb_dist,check", [(10.0, 100, True), (3.0, 64, True), (10.0, 100, False)])
def test_oracle_search_by_projection_kf_vs_python(proj_s, th, orb_dist, check):
    s = proj_s
    n, fm = O.search_by_projection_kf(s["kps"], s["desc"], s["occupied"], scenes.BOUNDS, scenes.SF, s["proj_xy"], s["level"], s["angle"],
                                      s["mp_desc"], th, orb_dist, check)
    pn, pfm = py_search_by_projection_kf(s, th, orb_dist, check)
    assert n == pn and np.array_equal(fm, pfm)
    assert n > 50


@pytest.mark.parametrize("only_stereo,check", [(False, True), (True, True), (False, False)])
def test_oracle_search_for_triangulation_vs_python(only_stereo, check):
    s = scenes.triangulation_scene(1000)
    n, mt = O.search_for_triangulation(*_tri_args(s), only_stereo, check)
    pn, pmt = py_search_for_triangulation(s, only_stereo, check)
    assert n == pn and np.array_equal(mt, pmt)
    assert n > (10 if only_stereo else 40)


# ------------------------------------------------------------------------------------------------ GPU: C ABI vs oracle
@pytest.mark.gpu
@pytest.mark.parametrize("seed", [1000, 1001])
@pytest.mark.parametrize("th,chi2", [(3.0, True), (4.0, False), (2.5, True)])
def test_cuda_fuse(seed, th, chi2):
    import orbx
    s = scenes.fuse_scene(seed)
    m = orbx.ORBmatcher(0.6, True)
    n, bi = m.Fuse(*_fuse_args(s), th, chi2)
    on, obi = O.fuse(*_fuse_args(s), th, chi2)
    assert n == on and np.array_equal(bi, obi)
    assert n > 50 and m.last_launches() == 1


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [1000, 1001])
def test_cuda_search_in_area_raw(seed):
    """The generic op with every optional input: static skip list, chi-square gate, skipped queries."""
    import orbx
    s = scenes.fuse_scene(seed)
    m = orbx.ORBmatcher(0.6, True)
    lvl = s["level"].copy()
    lvl[::7] = -1
    qr = (np.float32(3.0) * scenes.SF[np.maximum(lvl, 0)]).astype(np.float32)
    for kw in [dict(), dict(skip=s["skip"]), dict(u_right=s["u_right"], qur=s["proj_ur"], inv_level_sigma2=scenes.INV_SIGMA2, skip=s["skip"])]:
        bi, bd = m.SearchInArea(s["kps"], s["desc"], scenes.BOUNDS, s["proj_xy"], qr, lvl, s["mp_desc"], **kw)
        obi, obd = O.search_in_area(s["kps"], s["desc"], scenes.BOUNDS, s["proj_xy"], qr, lvl, s["mp_desc"], **kw)
        assert np.array_equal(bi, obi) and np.array_equal(bd, obd)
        assert (bi[::7] == -1).all() and (bi >= 0).sum() > 100


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [1000, 1001])
@pytest.mark.parametrize("th", [7.5, 3.0])
def test_cuda_search_by_sim3(seed, th):
    import orbx
    s = scenes.sim3_scene(seed)
    m = orbx.ORBmatcher(0.75, True)
    n, mt = m.SearchBySim3(*_sim3_args(s), th)
    on, omt = O.search_by_sim3(*_sim3_args(s), th)
    assert n == on and np.array_equal(mt, omt)
    assert n > 50


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [1000, 1001])
@pytest.mark.parametrize("th", [10.0, 4.0])
def test_cuda_search_by_projection_sim3(seed, th):
    import orbx
    s = scenes.projection_scene(seed)
    m = orbx.ORBmatcher(0.75, True)
    n, fm = m.SearchByProjectionSim3(s["kps"], s["desc"], s["occupied"], scenes.BOUNDS, scenes.SF, s["proj_xy"], s["level"], s["mp_desc"], th)
    on, ofm = O.search_by_projection_sim3(s["kps"], s["desc"], s["occupied"], scenes.BOUNDS, scenes.SF, s["proj_xy"], s["level"], s["mp_desc"], th)
    assert n == on and np.array_equal(fm, ofm)
    assert n > 50


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [1000, 1001])
@pytest.mark.parametrize("th,orb_dist,check", [(10.0, 100, True), (3.0, 64, True), (10.0, 100, False)])
def test_cuda_search_by_projection_kf(seed, th, orb_dist, check):
    import orbx
    s = scenes.projection_scene(seed)
    m = orbx.ORBmatcher(0.75, check)
    n, fm = m.SearchByProjectionKF(s["kps"], s["desc"], s["occupied"], scenes.BOUNDS, scenes.SF, s["proj_xy"], s["level"], s["angle"], s["mp_desc"],
                                   th, orb_dist)
    on, ofm = O.search_by_projection_kf(s["kps"], s["desc"], s["occupied"], scenes.BOUNDS, scenes.SF, s["proj_xy"], s["level"], s["angle"],
                                        s["mp_desc"], th, orb_dist, check)
    assert n == on and np.array_equal(fm, ofm)
    assert n > 50


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [1000, 1001])
@pytest.mark.parametrize("only_stereo,check", [(False, True), (True, True), (False, False)])
def test_cuda_search_for_triangulation(seed, only_stereo, check):
    import orbx
    s = scenes.triangulation_scene(seed)
    m = orbx.ORBmatcher(0.6, check)
    n, mt = m.SearchForTriangulation(*_tri_args(s), only_stereo)
    on, omt = O.search_for_triangulation(*_tri_args(s), only_stereo, check)
    assert n == on and np.array_equal(mt, omt)
    assert n > (10 if only_stereo else 40)


@pytest.mark.gpu
@pytest.mark.parametrize("check", [True, False])
def test_cuda_search_for_triangulation_many_survivors(check):
    """The device writes back at most 16 gate-passing candidates per query and falls back to the full distance list when a
    query has more: 40 copies of one feature (same descriptor, same position) in the second key frame force that path."""
    import orbx
    s = dict(scenes.triangulation_scene(1002))
    k2, d2 = s["k2"].copy(), s["d2"].copy()
    has2, ur2 = s["has_mp2"].copy(), s["u_right2"].copy()
    # a correspondence the plain scene matches: its copies overwrite 40 other rows of key frame 2
    n0, mt0 = O.search_for_triangulation(*_tri_args(s), False, check)
    f1 = int(np.flatnonzero(mt0 >= 0)[5])
    f2 = int(mt0[f1])
    rows = [j for j in range(len(k2)) if j != f2][100:140]
    for j in rows:
        k2[j], d2[j], has2[j], ur2[j] = k2[f2], d2[f2], 0, ur2[f2]
    s.update(k2=k2, d2=d2, has_mp2=has2, u_right2=ur2, fv2=scenes.feature_vector(d2, 16))
    m = orbx.ORBmatcher(0.6, check)
    n, mt = m.SearchForTriangulation(*_tri_args(s), False)
    on, omt = O.search_for_triangulation(*_tri_args(s), False, check)
    assert n == on and np.array_equal(mt, omt)
    assert m.last_launches() == 2 and n > 40          # the fallback ran


@pytest.mark.gpu
def test_cuda_map_searches_empty_inputs():
    import orbx
    s = scenes.fuse_scene(1000)
    m = orbx.ORBmatcher(0.6, True)
    e = lambda dt, *shape: np.zeros(shape, dt)
    n, bi = m.Fuse(s["kps"], s["desc"], s["u_right"], scenes.BOUNDS, scenes.SF, scenes.INV_SIGMA2, e(np.float32, 0, 2), e(np.float32, 0),
                   e(np.int32, 0), e(np.uint8, 0, 32))
    assert n == 0 and len(bi) == 0
    n, bi = m.Fuse(s["kps"][:0], s["desc"][:0], s["u_right"][:0], scenes.BOUNDS, scenes.SF, scenes.INV_SIGMA2, s["proj_xy"], s["proj_ur"],
                   s["level"], s["mp_desc"])
    assert n == 0 and (bi == -1).all()
    with pytest.raises(orbx.OrbxError):
        m.Fuse(*_fuse_args(s)[:8], np.full_like(s["level"], 9), s["mp_desc"])


# ------------------------------------------------------------------------------------------------ ComputeDistinctiveDescriptors
def _observation_sets(seed, sizes):
    """Descriptor sets like the observations of a MapPoint: one base descriptor seen again with a few flipped bits, sometimes
    twice the same row (ties), sometimes unrelated."""
    rng = np.random.default_rng(seed)
    rows, off = [], [0]
    for n in sizes:
        base = rng.integers(0, 256, 32).astype(np.uint8)
        obs = np.repeat(base[None], n, 0)
        obs = scenes.flip_bits(obs, rng, 40) if n else obs
        if n > 3:
            obs[rng.integers(0, n)] = obs[rng.integers(0, n)]
            obs[rng.integers(0, n)] = rng.integers(0, 256, 32)
        rows.append(obs)
        off.append(off[-1] + n)
    return np.concatenate(rows).astype(np.uint8).reshape(-1, 32), np.asarray(off, np.int32)


def py_distinctive(desc):
    n = len(desc)
    if n == 0:
        return -1
    D = np.array([[_dist(desc[i], desc[j]) for j in range(n)] for i in range(n)])
    med = np.sort(D, axis=1)[:, int(0.5 * (n - 1))]
    return int(np.argmin(med))          # first minimum, like `median < BestMedian`


SIZES = [0, 1, 2, 3, 4, 5, 8, 17, 32, 33, 64, 100] + [3, 6, 9] * 10


def test_oracle_distinctive_descriptors_vs_python():
    desc, off = _observation_sets(5, SIZES)
    best = O.distinctive_descriptors(desc, off)
    assert [int(b) for b in best] == [py_distinctive(desc[off[p]:off[p + 1]]) for p in range(len(off) - 1)]
    assert best[0] == -1 and best[1] == 0 and len(set(best.tolist())) > 5


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [5, 6])
def test_cuda_distinctive_descriptors(seed):
    import orbx
    rng = np.random.default_rng(seed)
    sizes = SIZES + [300, 700] + [int(v) for v in rng.integers(0, 40, 3000)]
    desc, off = _observation_sets(seed, sizes)
    m = orbx.ORBmatcher(0.6, True)
    best = m.ComputeDistinctiveDescriptors(desc, off)
    assert np.array_equal(best, O.distinctive_descriptors(desc, off))
    assert m.last_launches() == 1
    assert len(m.ComputeDistinctiveDescriptors(desc[:0], np.zeros(1, np.int32))) == 0
