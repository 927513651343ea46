"""orbm_frame (include/orbx.h): the Frame that lives on the device.  Its grid is the reference's AssignFeaturesToGrid (Frame.cc:230-245,
397-407) built by a counting sort on the device; the searches that take the handle must give, index for index, what the oracle's
restatements of SearchByProjection(Frame, MapPoints) (ORBmatcher.cc:45-129), SearchByProjection(Current, Last) (:1330-1472) and
SearchForInitialization (:405-520) give on the host arrays."""
import numpy as np
import pytest

import oracle_py as O
import orbx
from orbx import synth
import scenes

pytestmark = pytest.mark.gpu
BOUNDS = (0.0, 640.0, 0.0, 480.0)


def _host_grid(kps, bounds):
    """AssignFeaturesToGrid in numpy: column-major cells, insertion order inside a cell."""
    minX, maxX, minY, maxY = [np.float32(b) for b in bounds]
    invW, invH = np.float32(64) / (maxX - minX), np.float32(48) / (maxY - minY)
    rnd = lambda v: np.where(v >= 0, np.floor(v + np.float32(0.5)), -np.floor(-v + np.float32(0.5))).astype(np.int64)    # roundf
    px, py = rnd((kps["x"] - minX) * invW), rnd((kps["y"] - minY) * invH)
    ok = (px >= 0) & (px < 64) & (py >= 0) & (py < 48)
    cell = np.where(ok, px * 48 + py, 1 << 20)
    order = np.argsort(cell, kind="stable")
    order = order[ok[order]]
    start = np.searchsorted(cell[order], np.arange(64 * 48 + 1))
    return start.astype(np.int32), order.astype(np.int32)


@pytest.mark.parametrize("seed,bounds", [(1000, BOUNDS), (1001, (-13.5, 655.25, -9.0, 470.0)), (1002, (100.0, 400.0, 50.0, 300.0))])
def test_device_grid_is_assign_features_to_grid(seed, bounds):
    s = scenes.projection_scene(seed)
    m = orbx.ORBmatcher(0.8, True)
    F = orbx.DeviceFrame(m, s["kps"], s["desc"], bounds)
    cs, en = F.grid()
    hs, he = _host_grid(s["kps"], bounds)
    assert np.array_equal(cs, hs)
    assert np.array_equal(en[:hs[-1]], he)
    assert F.keypoints().tobytes() == np.ascontiguousarray(s["kps"]).tobytes()


@pytest.mark.parametrize("seed", [1000, 1001])
@pytest.mark.parametrize("th,nnratio", [(1.0, 0.8), (3.0, 0.8), (5.0, 0.6)])
def test_search_by_projection_on_a_device_frame(seed, th, nnratio):
    s = scenes.projection_scene(seed)
    m = orbx.ORBmatcher(nnratio, True)
    F = orbx.DeviceFrame(m, s["kps"], s["desc"], scenes.BOUNDS)
    for _ in range(2):          # the handle is reused across searches, as across the searches of one Tracking step
        n, fm = m.SearchByProjectionDev(F, s["u_right"], s["occupied"], scenes.SF, s["proj_xy"], s["proj_xr"], s["level"], s["view_cos"], s["mp_desc"],
                                        s["mp_has_obs"], th)
        on, ofm = O.search_by_projection(s["kps"], s["desc"], s["u_right"], s["occupied"], scenes.BOUNDS, scenes.SF, s["proj_xy"], s["proj_xr"],
                                         s["level"], s["view_cos"], s["mp_desc"], s["mp_has_obs"], th, nnratio)
        assert n == on and np.array_equal(fm, ofm)
    assert n > 50


@pytest.mark.parametrize("th,mode,check", [(15.0, 0, True), (7.0, 1, True), (15.0, 2, False), (30.0, 0, True)])
def test_search_by_projection_frame_on_a_device_frame(th, mode, check):
    s = scenes.projection_scene(1000)
    m = orbx.ORBmatcher(0.9, check)
    F = orbx.DeviceFrame(m, s["kps"], s["desc"], scenes.BOUNDS)
    n, fm = m.SearchByProjectionFrameDev(F, s["u_right"], s["occupied"], scenes.SF, s["proj_xy"], s["proj_xr"], s["level"], s["angle"], s["mp_desc"],
                                         s["mp_has_obs"], th, mode)
    on, ofm = O.search_by_projection_frame(s["kps"], s["desc"], s["u_right"], s["occupied"], scenes.BOUNDS, scenes.SF, s["proj_xy"], s["proj_xr"],
                                           s["level"], s["angle"], s["mp_desc"], s["mp_has_obs"], th, mode, check)
    assert n == on and np.array_equal(fm, ofm)
    assert n > 50


@pytest.mark.parametrize("seed,shift", [(1000, 3.0), (1001, 11.0)])
def test_search_for_initialization_on_device_frames_from_the_extractor(seed, shift):
    """Both frames come straight from the extractor's device results (no descriptor round trip); windows of 100 px hold more candidates
    than the 64 slots a query gets at first, so this also takes the retry with more slots."""
    ex1, ex2 = orbx.ORBextractor(2000, 1.2, 8, 20, 7), orbx.ORBextractor(2000, 1.2, 8, 20, 7)
    img1 = synth.rects(640, 480, seed)
    rng = np.random.default_rng(seed)
    img2 = np.clip(np.roll(img1, int(shift), axis=1).astype(np.int16) + rng.integers(-3, 4, img1.shape), 0, 255).astype(np.uint8)
    k1, d1 = ex1(img1)
    k2, d2 = ex2(img2)
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    for ratio, ori, win in [(0.9, True, 100), (0.7, False, 100), (0.9, True, 10), (0.9, True, 300)]:
        m = orbx.ORBmatcher(ratio, ori)
        F1 = orbx.DeviceFrame(m, bounds=BOUNDS, extractor=ex1)
        F2 = orbx.DeviceFrame(m, bounds=BOUNDS, extractor=ex2)
        assert F1.keypoints().tobytes() == k1.tobytes() and len(F2) == len(k2)
        n, m12, newprev = m.SearchForInitializationDev(F1, F2, prev, win)
        on, om12, oprev = O.search_for_initialization(k1, d1, k2, d2, BOUNDS, prev, win, ratio, ori)
        assert n == on and np.array_equal(m12, om12) and newprev.tobytes() == oprev.tobytes()
    assert on > 50


def test_device_frame_undistorts_like_frame_undistort_keypoints():
    ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7)
    k, d = ex(synth.rects(640, 480, 1003))
    m = orbx.ORBmatcher(0.8, True)
    K4, dist = (517.3, 516.5, 318.6, 255.3), np.array([0.2624, -0.9531, -0.0054, 0.0026, 1.1633], np.float32)      # TUM1.yaml
    want = O.undistort_keypoints(k, K4, dist)
    bounds = (-20.0, 660.0, -20.0, 500.0)
    F = orbx.DeviceFrame(m, bounds=bounds, extractor=ex, K4=K4, dist_coef=dist)
    got = F.keypoints()
    assert got.tobytes() == want.tobytes()
    cs, en = F.grid()
    hs, he = _host_grid(want, bounds)
    assert np.array_equal(cs, hs) and np.array_equal(en[:hs[-1]], he)
    # k1 == 0: keypoints stay as extracted (Frame.cc:421-425)
    F0 = orbx.DeviceFrame(m, bounds=bounds, extractor=ex, K4=K4, dist_coef=np.zeros(5, np.float32))
    assert F0.keypoints().tobytes() == k.tobytes()


def test_device_frame_edge_cases():
    m = orbx.ORBmatcher(0.8, True)
    E = orbx.DeviceFrame(m, np.zeros(0, orbx.KP_DTYPE), np.zeros((0, 32), np.uint8), BOUNDS)
    assert len(E) == 0
    s = scenes.projection_scene(1000)
    n, fm = m.SearchByProjectionDev(E, np.zeros(0, np.float32), np.zeros(0, np.uint8), scenes.SF, s["proj_xy"], s["proj_xr"], s["level"], s["view_cos"],
                                    s["mp_desc"], s["mp_has_obs"], 3.0)
    assert n == 0 and len(fm) == 0
    F = orbx.DeviceFrame(m, s["kps"], s["desc"], scenes.BOUNDS)
    n, fm = m.SearchByProjectionDev(F, s["u_right"], s["occupied"], scenes.SF, np.zeros((0, 2), np.float32), np.zeros(0, np.float32), np.zeros(0, np.int32),
                                    np.zeros(0, np.float32), np.zeros((0, 32), np.uint8), np.zeros(0, np.uint8), 3.0)
    assert n == 0 and (fm == -1).all()
    with pytest.raises(orbx.OrbxError):
        orbx.DeviceFrame(m, np.zeros(9000, orbx.KP_DTYPE), np.zeros((9000, 32), np.uint8), BOUNDS)
