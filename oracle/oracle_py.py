"""ctypes binding of the CPU ORACLE (oracle/liborb_oracle.so).

TEST INFRASTRUCTURE, NOT PRODUCT: only tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py may import this module.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liborb_oracle.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28


def build(force=False):
    if force or not os.path.exists(_LIB_PATH):
        subprocess.check_call(["make", "-C", _HERE, "liborb_oracle.so"] + (["-B"] if force else []))
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        vp, i, f, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
        L.orbo_create.restype = vp
        L.orbo_create.argtypes = [i, f, i, i, i]
        L.orbo_destroy.argtypes = [vp]
        L.orbo_set_modes.argtypes = [vp, i, i]
        L.orbo_extract.restype = i
        L.orbo_extract.argtypes = [vp, vp, i, i, sz, vp, vp, i]
        L.orbo_scale_tables.argtypes = [vp, vp, vp, vp, vp]
        L.orbo_features_per_level.argtypes = [vp, vp]
        L.orbo_umax.argtypes = [vp, vp]
        L.orbo_level_size.argtypes = [vp, i, vp, vp]
        L.orbo_level_padded.argtypes = [vp, i, vp]
        L.orbo_level_blurred.argtypes = [vp, i, vp]
        L.orbo_level_candidates.restype = i
        L.orbo_level_candidates.argtypes = [vp, i, vp, i]
        L.orbo_level_nkeys.restype = i
        L.orbo_level_nkeys.argtypes = [vp, i]
        L.orbo_stage_seconds.argtypes = [vp, vp]
        L.orbo_resize_linear_u8.argtypes = [vp, i, i, sz, vp, i, i, sz]
        L.orbo_border_reflect101.argtypes = [vp, i, i, sz, vp, i, sz]
        L.orbo_gauss7.argtypes = [vp, i, i, sz, vp, sz]
        L.orbo_fast.restype = i
        L.orbo_fast.argtypes = [vp, i, i, sz, i, i, vp, i]
        L.orbo_fast_atan2.restype = f
        L.orbo_fast_atan2.argtypes = [f, f]
        L.orbo_distribute.restype = i
        L.orbo_distribute.argtypes = [vp, i, i, i, i, i, i, i, vp, i]
        L.orbo_descriptor_distance.restype = i
        L.orbo_descriptor_distance.argtypes = [vp, vp]
        L.orbo_match_bruteforce.argtypes = [vp, i, vp, i, vp, vp, vp]
        L.orbo_match_candidates.argtypes = [vp, i, vp, i, vp, vp, vp, vp, vp, vp]
        L.orbo_grid_build.restype = vp
        L.orbo_grid_build.argtypes = [vp, i, f, f, f, f]
        L.orbo_grid_destroy.argtypes = [vp]
        L.orbo_grid_query.restype = i
        L.orbo_grid_query.argtypes = [vp, vp, f, f, f, i, i, vp, i]
        L.orbo_search_for_initialization.restype = i
        L.orbo_search_for_initialization.argtypes = [vp, vp, i, vp, vp, i, f, f, f, f, vp, vp, i, f, i]
        L.orbo_search_by_projection.restype = i
        L.orbo_search_by_projection.argtypes = [vp, vp, vp, vp, i, f, f, f, f, vp, i, vp, vp, vp, vp, vp, vp, i, f, f, vp]
        L.orbo_search_by_projection_frame.restype = i
        L.orbo_search_by_projection_frame.argtypes = [vp, vp, vp, vp, i, f, f, f, f, vp, i, vp, vp, vp, vp, vp, vp, i, f, i, i, vp]
        L.orbo_search_by_bow.restype = i
        L.orbo_search_by_bow.argtypes = [i, vp, vp, i, vp, vp, vp, vp, i, vp, vp, i, vp, vp, vp, vp, i, f, i, vp]
        L.orbo_search_in_area.argtypes = [vp, vp, vp, vp, i, f, f, f, f, vp, vp, vp, vp, vp, i, vp, vp, vp]
        L.orbo_fuse.restype = i
        L.orbo_fuse.argtypes = [vp, vp, vp, i, f, f, f, f, vp, vp, i, vp, vp, vp, vp, i, f, i, vp]
        L.orbo_search_by_sim3.restype = i
        L.orbo_search_by_sim3.argtypes = [vp, vp, i, vp, vp, i, f, f, f, f, vp, i, vp, vp, vp, vp, vp, vp, f, vp]
        L.orbo_search_by_projection_sim3.restype = i
        L.orbo_search_by_projection_sim3.argtypes = [vp, vp, vp, i, f, f, f, f, vp, i, vp, vp, vp, i, f, vp]
        L.orbo_search_by_projection_kf.restype = i
        L.orbo_search_by_projection_kf.argtypes = [vp, vp, vp, i, f, f, f, f, vp, i, vp, vp, vp, vp, i, f, i, i, vp]
        L.orbo_search_for_triangulation.restype = i
        L.orbo_search_for_triangulation.argtypes = [vp, vp, i, vp, vp, vp, vp, vp, i, vp, vp, i, vp, vp, vp, vp, vp, i, vp, f, f, vp, vp, i, i, i, vp]
        L.orbo_distinctive_descriptor.restype = i
        L.orbo_distinctive_descriptor.argtypes = [vp, i]
        L.orbo_distinctive_descriptors.argtypes = [vp, vp, i, vp]
        L.orbo_cvt_gray.argtypes = [vp, i, i, sz, i, i, vp, sz]
        L.orbo_undistort_points.argtypes = [vp, i, f, f, f, f, vp, i, vp]
        L.orbo_undistort_keypoints.argtypes = [vp, i, f, f, f, f, vp, i, vp]
        L.orbo_image_bounds.argtypes = [i, i, f, f, f, f, vp, i, vp]
        L.orbo_stereo_from_rgbd.argtypes = [vp, vp, i, vp, sz, f, vp, vp]
        L.orbo_voc_create.restype = vp
        L.orbo_voc_create.argtypes = [i, i, i, i, vp, vp, vp, i]
        L.orbo_voc_destroy.argtypes = [vp]
        L.orbo_voc_transform_raw.argtypes = [vp, vp, i, i, vp, vp, vp]
        L.orbo_voc_transform.restype = i
        L.orbo_voc_transform.argtypes = [vp, vp, i, i, vp, vp, vp, vp, vp, vp]
        L.orbo_stereo_matches.restype = i
        L.orbo_stereo_matches.argtypes = [vp, vp, i, vp, vp, i, vp, vp, vp, vp, vp, vp, vp, vp, i, f, f, f, vp, vp]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _u8(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    return a


class OracleExtractor:
    """Mirror of ORB_SLAM2::ORBextractor (include/ORBextractor.h:45-111) over the CPU oracle."""

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7, tie_mode=0, fma_mode=0):
        self.L = lib()
        self.nlevels = nlevels
        self.nfeatures = nfeatures
        self.h = self.L.orbo_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
        self.L.orbo_set_modes(self.h, tie_mode, fma_mode)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orbo_destroy(self.h)
            self.h = None

    def set_modes(self, tie_mode=0, fma_mode=0):
        self.L.orbo_set_modes(self.h, tie_mode, fma_mode)

    def __call__(self, img):
        img = _u8(img)
        hh, ww = img.shape
        cap = max(4 * self.nfeatures + 64 * self.nlevels, 1024)
        kps = np.zeros(cap, dtype=KP_DTYPE)
        desc = np.zeros((cap, 32), dtype=np.uint8)
        n = self.L.orbo_extract(self.h, _p(img), ww, hh, img.strides[0], _p(kps), _p(desc), cap)
        assert n <= cap
        return kps[:n].copy(), desc[:n].copy()

    def scale_tables(self):
        out = [np.zeros(self.nlevels, dtype=np.float32) for _ in range(4)]
        self.L.orbo_scale_tables(self.h, *[_p(o) for o in out])
        return out

    def features_per_level(self):
        o = np.zeros(self.nlevels, dtype=np.int32)
        self.L.orbo_features_per_level(self.h, _p(o))
        return o

    def umax(self):
        o = np.zeros(16, dtype=np.int32)
        self.L.orbo_umax(self.h, _p(o))
        return o

    def level_size(self, l):
        w, h = C.c_int(), C.c_int()
        self.L.orbo_level_size(self.h, l, C.byref(w), C.byref(h))
        return w.value, h.value

    def level_padded(self, l):
        w, h = self.level_size(l)
        o = np.zeros((h + 38, w + 38), dtype=np.uint8)
        self.L.orbo_level_padded(self.h, l, _p(o))
        return o

    def level_blurred(self, l):
        w, h = self.level_size(l)
        o = np.zeros((h, w), dtype=np.uint8)
        self.L.orbo_level_blurred(self.h, l, _p(o))
        return o

    def level_candidates(self, l):
        n = self.L.orbo_level_candidates(self.h, l, None, 0)
        o = np.zeros((max(n, 1), 3), dtype=np.int32)
        self.L.orbo_level_candidates(self.h, l, _p(o), n)
        return o[:n]

    def level_nkeys(self, l):
        return self.L.orbo_level_nkeys(self.h, l)

    def stage_seconds(self):
        o = np.zeros(6, dtype=np.float64)
        self.L.orbo_stage_seconds(self.h, _p(o))
        return o


def resize_linear(src, dw, dh):
    src = _u8(src)
    dst = np.zeros((dh, dw), dtype=np.uint8)
    lib().orbo_resize_linear_u8(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dw, dh, dw)
    return dst


def border_reflect101(src, b):
    src = _u8(src)
    h, w = src.shape
    dst = np.zeros((h + 2 * b, w + 2 * b), dtype=np.uint8)
    lib().orbo_border_reflect101(_p(src), w, h, src.strides[0], _p(dst), b, w + 2 * b)
    return dst


def gauss7(src):
    src = _u8(src)
    h, w = src.shape
    dst = np.zeros((h, w), dtype=np.uint8)
    lib().orbo_gauss7(_p(src), w, h, src.strides[0], _p(dst), w)
    return dst


def fast(roi, threshold, nms=True):
    """roi may be a non-contiguous 2-D view (row stride honoured)."""
    assert roi.dtype == np.uint8 and roi.strides[1] == 1
    h, w = roi.shape
    cap = w * h
    o = np.zeros((max(cap, 1), 3), dtype=np.int32)
    n = lib().orbo_fast(_p(roi), w, h, roi.strides[0], threshold, int(nms), _p(o), cap)
    return o[:n]


def fast_atan2(y, x):
    return lib().orbo_fast_atan2(float(y), float(x))


def distribute(xyr, minX, maxX, minY, maxY, N, tie_mode=0):
    xyr = np.ascontiguousarray(xyr, dtype=np.int32)
    n = len(xyr)
    out = np.zeros(max(n, 1), dtype=np.int32)
    m = lib().orbo_distribute(_p(xyr), n, minX, maxX, minY, maxY, N, tie_mode, _p(out), n)
    return out[:m]


def descriptor_distance(a, b):
    a, b = _u8(a), _u8(b)
    return lib().orbo_descriptor_distance(_p(a), _p(b))


def match_bruteforce(q, t):
    q, t = _u8(q), _u8(t)
    nq, nt = len(q), len(t)
    bi = np.zeros(nq, np.int32); bd = np.zeros(nq, np.int32); sd = np.zeros(nq, np.int32)
    lib().orbo_match_bruteforce(_p(q), nq, _p(t), nt, _p(bi), _p(bd), _p(sd))
    return bi, bd, sd


def match_candidates(q, t, off, idx):
    q, t = _u8(q), _u8(t)
    off = np.ascontiguousarray(off, np.int32); idx = np.ascontiguousarray(idx, np.int32)
    nq = len(q)
    bi = np.zeros(nq, np.int32); bd = np.zeros(nq, np.int32); si = np.zeros(nq, np.int32); sd = np.zeros(nq, np.int32)
    lib().orbo_match_candidates(_p(q), nq, _p(t), len(t), _p(off), _p(idx), _p(bi), _p(bd), _p(si), _p(sd))
    return bi, bd, si, sd


def grid_query_all(kps, queries, r, minLevel, maxLevel, bounds):
    """GetFeaturesInArea for every (x, y) in queries -> CSR (off, idx)."""
    kps = np.ascontiguousarray(kps)
    G = lib().orbo_grid_build(_p(kps), len(kps), *[float(b) for b in bounds])
    off = [0]
    idx = []
    buf = np.zeros(max(len(kps), 1), np.int32)
    for (x, y) in queries:
        n = lib().orbo_grid_query(G, _p(kps), float(x), float(y), float(r), minLevel, maxLevel, _p(buf), len(buf))
        idx.append(buf[:n].copy())
        off.append(off[-1] + n)
    lib().orbo_grid_destroy(G)
    return np.asarray(off, np.int32), (np.concatenate(idx) if idx else np.zeros(0, np.int32)).astype(np.int32)


def search_for_initialization(k1, d1, k2, d2, bounds, prev_matched, windowSize=100, nnratio=0.9, checkOri=True):
    k1 = np.ascontiguousarray(k1); k2 = np.ascontiguousarray(k2)
    d1, d2 = _u8(d1), _u8(d2)
    prev = np.ascontiguousarray(prev_matched, dtype=np.float32).copy()
    m12 = np.zeros(len(k1), np.int32)
    n = lib().orbo_search_for_initialization(_p(k1), _p(d1), len(k1), _p(k2), _p(d2), len(k2),
                                             *[float(b) for b in bounds], _p(prev), _p(m12), windowSize, nnratio, int(checkOri))
    return n, m12, prev


def _f32(a):
    return np.ascontiguousarray(a, np.float32)


def _pn(a):
    return None if a is None else _p(a)


def search_by_projection(kps, desc, u_right, occupied, bounds, sf, proj_xy, proj_xr, level, view_cos, mp_desc, mp_has_obs, th=1.0, nnratio=0.8):
    """ORBmatcher::SearchByProjection(Frame&, vpMapPoints, th) (ORBmatcher.cc:45-129) -> (nmatches, feature_match)."""
    kps = np.ascontiguousarray(kps)
    desc, mp_desc = _u8(desc), _u8(mp_desc)
    u_right, sf, proj_xy, proj_xr, view_cos = _f32(u_right), _f32(sf), _f32(proj_xy), _f32(proj_xr), _f32(view_cos)
    occupied, mp_has_obs, level = _u8(occupied), _u8(mp_has_obs), np.ascontiguousarray(level, np.int32)
    fm = np.zeros(max(len(kps), 1), np.int32)
    n = lib().orbo_search_by_projection(_p(kps), _p(desc), _p(u_right), _p(occupied), len(kps), *[float(b) for b in bounds], _p(sf), len(sf),
                                        _p(proj_xy), _p(proj_xr), _p(level), _p(view_cos), _p(mp_desc), _p(mp_has_obs), len(level),
                                        float(th), float(nnratio), _p(fm))
    return n, fm[:len(kps)]


def search_by_projection_frame(kps, desc, u_right, occupied, bounds, sf, proj_xy, proj_ur, last_octave, last_angle, mp_desc, mp_has_obs,
                               th, mode=0, checkOri=True):
    """ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (ORBmatcher.cc:1330-1472) -> (nmatches, feature_match)."""
    kps = np.ascontiguousarray(kps)
    desc, mp_desc = _u8(desc), _u8(mp_desc)
    u_right, sf, proj_xy, proj_ur, last_angle = _f32(u_right), _f32(sf), _f32(proj_xy), _f32(proj_ur), _f32(last_angle)
    occupied, mp_has_obs, last_octave = _u8(occupied), _u8(mp_has_obs), np.ascontiguousarray(last_octave, np.int32)
    fm = np.zeros(max(len(kps), 1), np.int32)
    n = lib().orbo_search_by_projection_frame(_p(kps), _p(desc), _p(u_right), _p(occupied), len(kps), *[float(b) for b in bounds], _p(sf), len(sf),
                                              _p(proj_xy), _p(proj_ur), _p(last_octave), _p(last_angle), _p(mp_desc), _p(mp_has_obs),
                                              len(last_octave), float(th), int(mode), int(checkOri), _p(fm))
    return n, fm[:len(kps)]


def search_by_bow(k1, d1, valid1, fv1, k2, d2, fv2, valid2=None, nnratio=0.7, checkOri=True):
    """ORBmatcher::SearchByBoW, key frame -> frame (ORBmatcher.cc:159-288) when valid2 is None, else key frame -> key frame
    (:522-655).  fv = (node ids, offsets, feature indices) -> (nmatches, matches)."""
    k1, k2 = np.ascontiguousarray(k1), np.ascontiguousarray(k2)
    d1, d2 = _u8(d1), _u8(d2)
    n1_, o1_, i1_ = [np.ascontiguousarray(a, np.int32) for a in fv1]
    n2_, o2_, i2_ = [np.ascontiguousarray(a, np.int32) for a in fv2]
    variant = 0 if valid2 is None else 1
    valid1 = _u8(valid1)
    v2 = None if valid2 is None else _u8(valid2)
    nout = len(k2) if variant == 0 else len(k1)
    out = np.zeros(max(nout, 1), np.int32)
    n = lib().orbo_search_by_bow(variant, _p(k1), _p(d1), len(k1), _p(valid1), _p(n1_), _p(o1_), _p(i1_), len(n1_), _p(k2), _p(d2), len(k2),
                                 _pn(v2), _p(n2_), _p(o2_), _p(i2_), len(n2_), float(nnratio), int(checkOri), _p(out))
    return n, out[:nout]


def _i32(a):
    return np.ascontiguousarray(a, np.int32)


def search_in_area(kps, desc, bounds, qxy, qr, qlevel, qdesc, u_right=None, skip=None, qur=None, inv_level_sigma2=None):
    """The window + level + (optional chi-square) + argmin loop shared by Fuse and SearchBySim3 (ORBmatcher.cc:886-935,
    :1059-1077, :1183-1203) -> (best_idx, best_dist)."""
    kps = np.ascontiguousarray(kps)
    desc, qdesc = _u8(desc), _u8(qdesc)
    qxy, qr, qlevel = _f32(qxy), _f32(qr), _i32(qlevel)
    ur = None if u_right is None else _f32(u_right)
    sk = None if skip is None else _u8(skip)
    qu = None if qur is None else _f32(qur)
    inv = None if inv_level_sigma2 is None else _f32(inv_level_sigma2)
    nq = len(qlevel)
    bi, bd = np.zeros(max(nq, 1), np.int32), np.zeros(max(nq, 1), np.int32)
    lib().orbo_search_in_area(_p(kps), _p(desc), _pn(ur), _pn(sk), len(kps), *[float(b) for b in bounds], _p(qxy), _pn(qu), _p(qr),
                              _p(qlevel), _p(qdesc), nq, _pn(inv), _p(bi), _p(bd))
    return bi[:nq], bd[:nq]


def fuse(kps, desc, u_right, bounds, sf, inv_level_sigma2, proj_xy, proj_ur, level, mp_desc, th=3.0, stereo_chi2=True):
    """ORBmatcher::Fuse(KF, MapPoints, th) (ORBmatcher.cc:827-975) / Fuse(KF, Scw, ...) (:977-1102, stereo_chi2=False)
    -> (nFused, best_idx per point)."""
    kps = np.ascontiguousarray(kps)
    desc, mp_desc = _u8(desc), _u8(mp_desc)
    u_right, sf, inv, proj_xy, proj_ur, level = _f32(u_right), _f32(sf), _f32(inv_level_sigma2), _f32(proj_xy), _f32(proj_ur), _i32(level)
    out = np.zeros(max(len(level), 1), np.int32)
    n = lib().orbo_fuse(_p(kps), _p(desc), _p(u_right), len(kps), *[float(b) for b in bounds], _p(sf), _p(inv), len(sf), _p(proj_xy),
                        _p(proj_ur), _p(level), _p(mp_desc), len(level), float(th), int(stereo_chi2), _p(out))
    return n, out[:len(level)]


def search_by_sim3(k1, d1, k2, d2, bounds, sf, proj12_xy, level12, mp1_desc, proj21_xy, level21, mp2_desc, th=7.5):
    """ORBmatcher::SearchBySim3 (ORBmatcher.cc:1104-1328) -> (nFound, match12)."""
    k1, k2 = np.ascontiguousarray(k1), np.ascontiguousarray(k2)
    d1, d2, m1, m2 = _u8(d1), _u8(d2), _u8(mp1_desc), _u8(mp2_desc)
    sf, p12, p21, l12, l21 = _f32(sf), _f32(proj12_xy), _f32(proj21_xy), _i32(level12), _i32(level21)
    out = np.zeros(max(len(k1), 1), np.int32)
    n = lib().orbo_search_by_sim3(_p(k1), _p(d1), len(k1), _p(k2), _p(d2), len(k2), *[float(b) for b in bounds], _p(sf), len(sf),
                                  _p(p12), _p(l12), _p(m1), _p(p21), _p(l21), _p(m2), float(th), _p(out))
    return n, out[:len(k1)]


def search_by_projection_sim3(kps, desc, matched, bounds, sf, proj_xy, level, mp_desc, th=10.0):
    """ORBmatcher::SearchByProjection(KF, Scw, vpPoints, vpMatched, th) (ORBmatcher.cc:290-403) -> (nmatches, feature_match)."""
    kps = np.ascontiguousarray(kps)
    desc, mp_desc, matched = _u8(desc), _u8(mp_desc), _u8(matched)
    sf, proj_xy, level = _f32(sf), _f32(proj_xy), _i32(level)
    fm = np.zeros(max(len(kps), 1), np.int32)
    n = lib().orbo_search_by_projection_sim3(_p(kps), _p(desc), _p(matched), len(kps), *[float(b) for b in bounds], _p(sf), len(sf),
                                             _p(proj_xy), _p(level), _p(mp_desc), len(level), float(th), _p(fm))
    return n, fm[:len(kps)]


def search_by_projection_kf(kps, desc, occupied, bounds, sf, proj_xy, level, kf_angle, mp_desc, th=10.0, orb_dist=100, checkOri=True):
    """ORBmatcher::SearchByProjection(CurrentFrame, KF, sAlreadyFound, th, ORBdist) (ORBmatcher.cc:1474-1601)."""
    kps = np.ascontiguousarray(kps)
    desc, mp_desc, occupied = _u8(desc), _u8(mp_desc), _u8(occupied)
    sf, proj_xy, level, kf_angle = _f32(sf), _f32(proj_xy), _i32(level), _f32(kf_angle)
    fm = np.zeros(max(len(kps), 1), np.int32)
    n = lib().orbo_search_by_projection_kf(_p(kps), _p(desc), _p(occupied), len(kps), *[float(b) for b in bounds], _p(sf), len(sf),
                                           _p(proj_xy), _p(level), _p(kf_angle), _p(mp_desc), len(level), float(th), int(orb_dist),
                                           int(checkOri), _p(fm))
    return n, fm[:len(kps)]


def search_for_triangulation(k1, d1, has_mp1, u_right1, fv1, k2, d2, has_mp2, u_right2, fv2, F12, epipole, sf2, sigma2_2,
                             only_stereo=False, checkOri=True):
    """ORBmatcher::SearchForTriangulation (ORBmatcher.cc:657-825) -> (nmatches, matches12)."""
    k1, k2 = np.ascontiguousarray(k1), np.ascontiguousarray(k2)
    d1, d2, h1, h2 = _u8(d1), _u8(d2), _u8(has_mp1), _u8(has_mp2)
    u1, u2, F, sf2, sg = _f32(u_right1), _f32(u_right2), _f32(np.asarray(F12).reshape(9)), _f32(sf2), _f32(sigma2_2)
    n1_, o1_, i1_ = [_i32(a) for a in fv1]
    n2_, o2_, i2_ = [_i32(a) for a in fv2]
    out = np.zeros(max(len(k1), 1), np.int32)
    n = lib().orbo_search_for_triangulation(_p(k1), _p(d1), len(k1), _p(h1), _p(u1), _p(n1_), _p(o1_), _p(i1_), len(n1_),
                                            _p(k2), _p(d2), len(k2), _p(h2), _p(u2), _p(n2_), _p(o2_), _p(i2_), len(n2_),
                                            _p(F), float(epipole[0]), float(epipole[1]), _p(sf2), _p(sg), len(sf2),
                                            int(only_stereo), int(checkOri), _p(out))
    return n, out[:len(k1)]


def cvt_gray(img, rgb=True):
    """cvtColor(img, CV_RGB2GRAY | CV_BGR2GRAY | CV_RGBA2GRAY | CV_BGRA2GRAY) (Tracking.cc:280-305); img: H x W x (3|4) uint8."""
    img = _u8(img)
    h, w, cn = img.shape
    out = np.zeros((h, w), np.uint8)
    lib().orbo_cvt_gray(_p(img), w, h, img.strides[0], cn, int(rgb), _p(out), w)
    return out


def undistort_points(pts, K4, dist):
    """cv::undistortPoints(pts, K, dist, noArray(), K) (Frame.cc:438); K4 = (fx, fy, cx, cy), dist = (k1, k2, p1, p2[, k3])."""
    pts, dist = _f32(pts).reshape(-1, 2), _f32(dist).reshape(-1)
    out = np.zeros_like(pts)
    lib().orbo_undistort_points(_p(pts), len(pts), *[float(np.float32(v)) for v in K4], _p(dist), len(dist), _p(out))
    return out


def undistort_keypoints(kps, K4, dist):
    """Frame::UndistortKeyPoints (Frame.cc:419-449) -> mvKeysUn."""
    kps, dist = np.ascontiguousarray(kps), _f32(dist).reshape(-1)
    out = np.zeros_like(kps)
    lib().orbo_undistort_keypoints(_p(kps), len(kps), *[float(np.float32(v)) for v in K4], _p(dist), len(dist), _p(out))
    return out


def image_bounds(cols, rows, K4, dist):
    """Frame::ComputeImageBounds (Frame.cc:451-479) -> (mnMinX, mnMaxX, mnMinY, mnMaxY)."""
    dist = _f32(dist).reshape(-1)
    out = np.zeros(4, np.float32)
    lib().orbo_image_bounds(cols, rows, *[float(np.float32(v)) for v in K4], _p(dist), len(dist), _p(out))
    return out


def stereo_from_rgbd(kps, kps_un, im_depth, bf):
    """Frame::ComputeStereoFromRGBD (Frame.cc:658-679) -> (mvuRight, mvDepth)."""
    kps, kps_un, im = np.ascontiguousarray(kps), np.ascontiguousarray(kps_un), _f32(im_depth)
    uR, dep = np.zeros(len(kps), np.float32), np.zeros(len(kps), np.float32)
    lib().orbo_stereo_from_rgbd(_p(kps), _p(kps_un), len(kps), _p(im), im.strides[0], float(bf), _p(uR), _p(dep))
    return uR, dep


class OracleVocabulary:
    """ORBVocabulary (DBoW2::TemplatedVocabulary<FORB>) over arrays: parent[i] of node i (node 0 = root), 32-byte node descriptors,
    node weights.  scoring / weighting use DBoW2's enum values (L1_NORM = 0, TF_IDF = 0)."""

    def __init__(self, k, L, parent, desc, weight, scoring=0, weighting=0):
        self.parent, self.desc, self.weight = _i32(parent), _u8(desc), np.ascontiguousarray(weight, np.float64)
        self.L = lib()
        self.h = self.L.orbo_voc_create(k, L, scoring, weighting, _p(self.parent), _p(self.desc), _p(self.weight), len(self.parent))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orbo_voc_destroy(self.h)
            self.h = None

    def transform_raw(self, desc, levelsup=4):
        desc = _u8(desc)
        n = len(desc)
        wid, nid, w = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.float64)
        self.L.orbo_voc_transform_raw(self.h, _p(desc), n, levelsup, _p(wid), _p(w), _p(nid))
        return wid[:n], w[:n], nid[:n]

    def transform(self, desc, levelsup=4):
        """-> (bow ids, bow values), (fv nodes, fv offsets, fv feature indices)"""
        desc = _u8(desc)
        n = len(desc)
        bid, bval = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.float64)
        fn, fo, fi = np.zeros(max(n, 1), np.int32), np.zeros(n + 1, np.int32), np.zeros(max(n, 1), np.int32)
        nfv = C.c_int(0)
        nb = self.L.orbo_voc_transform(self.h, _p(desc), n, levelsup, _p(bid), _p(bval), _p(fn), _p(fo), _p(fi), C.byref(nfv))
        return (bid[:nb], bval[:nb]), (fn[:nfv.value], fo[:nfv.value + 1], fi[:fo[nfv.value]])


def distinctive_descriptors(desc, off):
    """MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:281-311) for many MapPoints: observations of point p are rows
    off[p]..off[p+1] of desc -> index (within the point's rows) of its most distinctive descriptor, -1 when it has none."""
    desc, off = _u8(desc), _i32(off)
    best = np.zeros(max(len(off) - 1, 1), np.int32)
    lib().orbo_distinctive_descriptors(_p(desc), _p(off), len(off) - 1, _p(best))
    return best[:len(off) - 1]


def stereo_matches(kl, dl, kr, dr, pyrL, pyrR, sf, isf, bf, maxD, minD=0.0):
    """pyrL/pyrR: lists of 2-D uint8 arrays (level interiors, any row stride)."""
    kl = np.ascontiguousarray(kl); kr = np.ascontiguousarray(kr)
    dl, dr = _u8(dl), _u8(dr)
    nl = len(pyrL)
    PL = (C.c_void_p * nl)(*[p.ctypes.data for p in pyrL])
    PR = (C.c_void_p * nl)(*[p.ctypes.data for p in pyrR])
    lw = np.array([p.shape[1] for p in pyrL], np.int32)
    lh = np.array([p.shape[0] for p in pyrL], np.int32)
    sL = np.array([p.strides[0] for p in pyrL], np.uint64)
    sR = np.array([p.strides[0] for p in pyrR], np.uint64)
    sf = np.ascontiguousarray(sf, np.float32); isf = np.ascontiguousarray(isf, np.float32)
    uR = np.zeros(len(kl), np.float32); dep = np.zeros(len(kl), np.float32)
    n = lib().orbo_stereo_matches(_p(kl), _p(dl), len(kl), _p(kr), _p(dr), len(kr), PL, PR, _p(lw), _p(lh), _p(sL), _p(sR),
                                  _p(sf), _p(isf), nl, bf, minD, maxD, _p(uR), _p(dep))
    return n, uR, dep
