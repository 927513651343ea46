"""orbx -- Python mirror of the reference's front-end interface over the C ABI of liborbx.so (include/orbx.h).

`ORBextractor` / `ORBmatcher` keep the reference's names, argument meaning and error behaviour
(include/ORBextractor.h:45-111, include/ORBmatcher.h:37-102 of kimwin2/ORB_SLAM2v2-1); every compute call goes
through ctypes into hand-written sm_100a CUDA.  There is no CPU fallback: importing works anywhere (so that the
ABI can be inspected), but creating an extractor or matcher without the library or without a CUDA device raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ORBX_LIB", os.path.join(_HERE, "liborbx.so"))      # ORBX_LIB: developer override (kernel A/B builds)

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28

OK, ERR_ARG, ERR_CUDA, ERR_CAPACITY, ERR_STATE = range(5)


class OrbxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("orbx error %d: %s" % (code, msg))
        self.code = code


_lib = None


def lib():
    """Load liborbx.so (built in-tree by __graft_entry__.build() / make -C csrc).  Fails loudly when absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise OrbxError(ERR_STATE, "liborbx.so is not built (%s): run `python -c 'import __graft_entry__ as g; g.build()'`; "
                        "there is no CPU fallback" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, i, f, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
    sig = {
        "orbx_create": (i, [i, f, i, i, i, i, i, vp]),
        "orbx_destroy": (None, [vp]),
        "orbx_alloc_pinned": (i, [sz, i, vp]),
        "orbx_free_pinned": (i, [vp]),
        "orbx_set_fma_mode": (i, [vp, i]),
        "orbx_set_octree_form": (i, [vp, i]),
        "orbx_levels": (i, [vp]),
        "orbx_scale_tables": (i, [vp, vp, vp, vp, vp]),
        "orbx_features_per_level": (i, [vp, vp]),
        "orbx_umax": (i, [vp, vp]),
        "orbx_max_keypoints": (i, [vp]),
        "orbx_extract": (i, [vp, vp, i, i, sz, vp, vp, i, vp]),
        "orbx_extract_batch": (i, [vp, vp, i, i, i, sz, vp, vp, i, vp]),
        "orbx_extract_device": (i, [vp, vp, i, i, i, sz, sz]),
        "orbx_fetch": (i, [vp, vp, vp, i, vp]),
        "orbx_pool_create": (i, [vp, i, i, f, i, i, i, i, i, vp]),
        "orbx_pool_destroy": (None, [vp]),
        "orbx_pool_devices": (i, [vp]),
        "orbx_pool_set_fma_mode": (i, [vp, i]),
        "orbx_pool_plan": (i, [vp, i, i, vp]),
        "orbx_pool_extract_batch": (i, [vp, vp, i, i, i, sz, vp, vp, i, vp]),
        "orbx_pool_submit_batch": (i, [vp, vp, i, i, i, sz, vp, vp, i, vp, vp]),
        "orbx_pool_wait": (i, [vp, C.c_ulonglong]),
        "orbx_results_device": (i, [vp, vp, vp, vp, vp]),
        "orbx_sync": (i, [vp]),
        "orbx_stream": (vp, [vp]),
        "orbx_last_launches": (i, [vp]),
        "orbx_set_profiling": (i, [vp, i]),
        "orbx_stage_ms": (i, [vp, vp]),
        "orbx_level_size": (i, [vp, i, vp, vp]),
        "orbx_pyramid_level": (i, [vp, i, i, i, vp, sz]),
        "orbx_pyramid_device": (i, [vp, i, i, vp, vp]),
        "orbx_set_pyramid_mirror": (i, [vp, i]),
        "orbx_pyramid_mirror": (i, [vp, i, vp, vp]),
        "orbx_debug_blurred_level": (i, [vp, i, i, vp, sz]),
        "orbx_debug_candidates": (i, [vp, i, i, vp, i, vp]),
        "orbx_debug_level_count": (i, [vp, i, i, vp]),
        "orbm_create": (i, [i, vp]),
        "orbm_destroy": (None, [vp]),
        "orbm_trim": (None, []),
        "orbm_frame_create": (i, [vp, vp, vp, i, f, f, f, f, vp]),
        "orbm_frame_from_extractor": (i, [vp, vp, i, f, f, f, f, vp, i, f, f, f, f, vp]),
        "orbm_frame_destroy": (None, [vp]),
        "orbm_frame_size": (i, [vp]),
        "orbm_frame_keypoints": (i, [vp, vp]),
        "orbm_frame_grid": (i, [vp, vp, vp, vp]),
        "orbm_search_by_projection_dev": (i, [vp, vp, vp, vp, vp, i, vp, vp, vp, vp, vp, vp, i, f, f, vp, vp]),
        "orbm_search_by_projection_frame_dev": (i, [vp, vp, vp, vp, vp, i, vp, vp, vp, vp, vp, vp, i, f, i, i, vp, vp]),
        "orbm_search_for_initialization_dev": (i, [vp, vp, vp, vp, vp, i, f, i, vp]),
        "orbm_device": (i, [vp]),
        "orbm_stream": (vp, [vp]),
        "orbm_last_launches": (i, [vp]),
        "orbm_descriptor_distance": (i, [vp, vp, vp, i, vp]),
        "orbm_match_bruteforce": (i, [vp, vp, i, vp, i, vp, vp, vp]),
        "orbm_match_ratio": (i, [vp, vp, i, vp, i, i, f, vp, vp, vp, vp]),
        "orbm_match_bruteforce_device": (i, [vp, vp, i, vp, i, vp, vp, vp]),
        "orbm_match_bruteforce_device_noalloc": (i, [vp, i, vp, i, vp, vp, vp, vp]),
        "orbm_match_candidates": (i, [vp, vp, i, vp, i, vp, vp, vp, vp, vp, vp]),
        "orbm_candidate_distances": (i, [vp, vp, i, vp, i, vp, vp, vp]),
        "orbm_grid_candidates": (i, [vp, vp, i, f, f, f, f, vp, i, vp, i, i, vp, vp, vp, i]),
        "orbm_search_for_initialization": (i, [vp, vp, vp, i, vp, vp, i, f, f, f, f, vp, vp, i, f, i, vp]),
        "orbm_stereo_matches": (i, [vp, vp, vp, i, vp, vp, i, vp, vp, i, f, f, f, vp, vp, vp]),
        "orbm_stereo_matches_batch": (i, [vp, vp, vp, f, f, f, vp, vp, i, vp]),
        "orbx_pyramid_layout": (i, [vp, i, vp, vp, vp]),
        "orbx_last_batch": (i, [vp]),
        "orbm_search_by_projection": (i, [vp, vp, vp, vp, vp, i, f, f, f, f, vp, i, vp, vp, vp, vp, vp, vp, i, f, f, vp, vp]),
        "orbm_search_by_projection_frame": (i, [vp, vp, vp, vp, vp, i, f, f, f, f, vp, i, vp, vp, vp, vp, vp, vp, i, f, i, i, vp, vp]),
        "orbm_search_by_bow": (i, [vp, i, vp, vp, i, vp, vp, vp, vp, i, vp, vp, i, vp, vp, vp, vp, i, f, i, vp, vp]),
        "orbm_search_in_area": (i, [vp, vp, vp, vp, vp, i, f, f, f, f, vp, vp, vp, vp, vp, i, vp, i, vp, vp]),
        "orbm_fuse": (i, [vp, vp, vp, vp, i, f, f, f, f, vp, vp, i, vp, vp, vp, vp, i, f, i, vp, vp]),
        "orbm_search_by_sim3": (i, [vp, vp, vp, i, vp, vp, i, f, f, f, f, vp, i, vp, vp, vp, vp, vp, vp, f, vp, vp]),
        "orbm_search_by_projection_sim3": (i, [vp, vp, vp, vp, i, f, f, f, f, vp, i, vp, vp, vp, i, f, vp, vp]),
        "orbm_search_by_projection_kf": (i, [vp, vp, vp, vp, i, f, f, f, f, vp, i, vp, vp, vp, vp, i, f, i, i, vp, vp]),
        "orbm_search_for_triangulation": (i, [vp, vp, vp, i, vp, vp, vp, vp, vp, i, vp, vp, i, vp, vp, vp, vp, vp, i, vp, f, f, vp, vp, i,
                                              i, i, vp, vp]),
        "orbm_distinctive_descriptors": (i, [vp, vp, vp, i, vp]),
        "orbx_extract_color_batch": (i, [vp, vp, i, i, i, C.c_size_t, i, i, vp, vp, i, vp]),
        "orbx_extract_color_device": (i, [vp, vp, i, i, i, C.c_size_t, C.c_size_t, i, i]),
        "orbm_undistort_points": (i, [vp, vp, i, f, f, f, f, vp, i, vp]),
        "orbm_undistort_keypoints": (i, [vp, vp, i, f, f, f, f, vp, i, vp]),
        "orbm_image_bounds": (i, [vp, i, i, f, f, f, f, vp, i, vp]),
        "orbm_stereo_from_rgbd": (i, [vp, vp, vp, i, vp, i, i, C.c_size_t, f, vp, vp]),
        "orbv_create": (i, [i, i, i, i, i, vp, vp, vp, i, vp]),
        "orbv_destroy": (None, [vp]),
        "orbv_size": (i, [vp]),
        "orbv_transform_raw": (i, [vp, vp, i, i, vp, vp, vp]),
        "orbv_transform": (i, [vp, vp, i, i, vp, vp, vp, vp, vp, vp, vp]),
        "orbx_last_error": (C.c_char_p, []),
        "orbx_version": (C.c_char_p, []),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


ABI_SYMBOLS = None  # filled lazily by abi_symbols()


def _check(rc, allow=()):
    if rc != OK and rc not in allow:
        raise OrbxError(rc, lib().orbx_last_error().decode())
    return rc


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _u8(a):
    return np.ascontiguousarray(a, dtype=np.uint8)


class ORBextractor:
    """ORB_SLAM2::ORBextractor (include/ORBextractor.h:45-111).

    ex = ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST); kps, desc = ex(image).
    `mask` is accepted and ignored like the reference ("Mask is ignored in the current implementation").
    An empty image returns (empty, empty) like the reference's silent return (src/ORBextractor.cc:1046-1047).
    """

    HARRIS_SCORE, FAST_SCORE = 0, 1

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7, device=0, max_batch=1):
        self._L = lib()
        h = C.c_void_p()
        _check(self._L.orbx_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device, max_batch, C.byref(h)))
        self._h = h
        self.nfeatures, self.nlevels, self.device = nfeatures, nlevels, device
        self.last_batch = 0

    def close(self):
        if getattr(self, "_h", None):
            self._L.orbx_destroy(self._h)
            self._h = None

    __del__ = close

    # --- accessors (ORBextractor.h:63-83)
    def GetLevels(self):
        return self._L.orbx_levels(self._h)

    def _tables(self):
        out = [np.zeros(self.nlevels, np.float32) for _ in range(4)]
        _check(self._L.orbx_scale_tables(self._h, *[_p(o) for o in out]))
        return out

    def GetScaleFactor(self):
        return float(self._tables()[0][1]) if self.nlevels > 1 else 1.0

    def GetScaleFactors(self):
        return self._tables()[0]

    def GetInverseScaleFactors(self):
        return self._tables()[1]

    def GetScaleSigmaSquares(self):
        return self._tables()[2]

    def GetInverseScaleSigmaSquares(self):
        return self._tables()[3]

    def features_per_level(self):
        o = np.zeros(self.nlevels, np.int32)
        _check(self._L.orbx_features_per_level(self._h, _p(o)))
        return o

    def umax(self):
        o = np.zeros(16, np.int32)
        _check(self._L.orbx_umax(self._h, _p(o)))
        return o

    def set_fma_mode(self, mode):
        _check(self._L.orbx_set_fma_mode(self._h, int(mode)))

    def set_octree_form(self, form):
        """0 = automatic, 1 / 2 / 3 = the 128- / 256- / 512-thread form of the quadtree kernel (identical results)."""
        _check(self._L.orbx_set_octree_form(self._h, int(form)))

    def max_keypoints(self):
        return self._L.orbx_max_keypoints(self._h)

    # --- operator()
    def __call__(self, image, mask=None):
        if image is None or image.size == 0:
            return np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        assert image.dtype == np.uint8 and image.ndim == 2, "CV_8UC1 expected"
        if image.strides[1] != 1:
            image = np.ascontiguousarray(image)
        hh, ww = image.shape
        cap = max(self.max_keypoints(), 64)
        for _ in range(2):
            kps = np.zeros(cap, KP_DTYPE)
            desc = np.zeros((cap, 32), np.uint8)
            n = C.c_int(0)
            rc = _check(self._L.orbx_extract(self._h, _p(image), ww, hh, image.strides[0], _p(kps), _p(desc), cap, C.byref(n)),
                        allow=(ERR_CAPACITY,))
            self.last_batch = 1
            if rc == OK:
                break
            cap = n.value
        return kps[:n.value].copy(), desc[:n.value].copy()

    def extract_batch(self, images):
        """images: sequence of equally-shaped 2-D uint8 arrays -> list of (kps, desc)."""
        imgs = [_u8(im) for im in images]
        B = len(imgs)
        hh, ww = imgs[0].shape
        assert all(im.shape == (hh, ww) for im in imgs)
        ptrs = (C.c_void_p * B)(*[im.ctypes.data for im in imgs])
        # the capacity depends on the shape: probe with a first call that may report ORBX_ERR_CAPACITY
        cap = max(self.max_keypoints(), 64)
        for _ in range(2):
            kps = np.zeros((B, cap), KP_DTYPE)
            desc = np.zeros((B, cap, 32), np.uint8)
            n = np.zeros(B, np.int32)
            rc = _check(self._L.orbx_extract_batch(self._h, ptrs, B, ww, hh, ww, _p(kps), _p(desc), cap, _p(n)), allow=(ERR_CAPACITY,))
            self.last_batch = B
            if rc == OK:
                break
            cap = int(n.max())
        return [(kps[b, :n[b]].copy(), desc[b, :n[b]].copy()) for b in range(B)]

    def extract_color_batch(self, images, rgb=True):
        """images: sequence of equally-shaped H x W x (3|4) uint8 arrays (RGB(A) when rgb, else BGR(A)): grey conversion
        (Tracking.cc:280-305) + extraction on the device -> list of (kps, desc)."""
        imgs = [_u8(im) for im in images]
        B = len(imgs)
        hh, ww, cn = imgs[0].shape
        assert all(im.shape == (hh, ww, cn) for im in imgs)
        ptrs = (C.c_void_p * B)(*[im.ctypes.data for im in imgs])
        cap = max(self.max_keypoints(), 64)
        for _ in range(2):
            kps = np.zeros((B, cap), KP_DTYPE)
            desc = np.zeros((B, cap, 32), np.uint8)
            n = np.zeros(B, np.int32)
            rc = _check(self._L.orbx_extract_color_batch(self._h, ptrs, B, ww, hh, ww * cn, cn, int(rgb), _p(kps), _p(desc), cap, _p(n)),
                        allow=(ERR_CAPACITY,))
            self.last_batch = B
            if rc == OK:
                break
            cap = int(n.max())
        return [(kps[b, :n[b]].copy(), desc[b, :n[b]].copy()) for b in range(B)]

    def extract_color_device(self, d_ptr, B, w, h, pitch, frame_stride, channels, rgb=True):
        """Device-resident colour frames: grey conversion into the handle's level-0 planes, then the extraction (no synchronisation)."""
        _check(self._L.orbx_extract_color_device(self._h, C.c_void_p(d_ptr), B, w, h, pitch, frame_stride, channels, int(rgb)))
        self.last_batch = B

    def extract_device(self, d_ptr, B, w, h, pitch, frame_stride):
        """Enqueue extraction of B device-resident frames on the handle's stream (no synchronisation)."""
        _check(self._L.orbx_extract_device(self._h, C.c_void_p(d_ptr), B, w, h, pitch, frame_stride))
        self.last_batch = B

    def fetch(self):
        B = self.last_batch
        cap = max(self.max_keypoints(), 64)
        kps = np.zeros((B, cap), KP_DTYPE)
        desc = np.zeros((B, cap, 32), np.uint8)
        n = np.zeros(B, np.int32)
        _check(self._L.orbx_fetch(self._h, _p(kps), _p(desc), cap, _p(n)))
        return [(kps[b, :n[b]].copy(), desc[b, :n[b]].copy()) for b in range(B)]

    def sync(self):
        _check(self._L.orbx_sync(self._h))

    def stream(self):
        return self._L.orbx_stream(self._h)

    def last_launches(self):
        return self._L.orbx_last_launches(self._h)

    STAGES = ("pyramid", "fast_cells", "octree", "blur", "orient_describe")

    def set_profiling(self, on=True):
        _check(self._L.orbx_set_profiling(self._h, int(on)))

    def stage_ms(self):
        o = np.zeros(5, np.float32)
        _check(self._L.orbx_stage_ms(self._h, _p(o)))
        return dict(zip(self.STAGES, [float(v) for v in o]))

    # --- mvImagePyramid (ORBextractor.h:85)
    def level_size(self, level):
        w, h = C.c_int(), C.c_int()
        _check(self._L.orbx_level_size(self._h, level, C.byref(w), C.byref(h)))
        return w.value, h.value

    def pyramid_level(self, level, frame=0, border=0):
        w, h = self.level_size(level)
        o = np.zeros((h + 2 * border, w + 2 * border), np.uint8)
        _check(self._L.orbx_pyramid_level(self._h, frame, level, border, _p(o), o.strides[0]))
        return o

    @property
    def mvImagePyramid(self):
        return [self.pyramid_level(l) for l in range(self.nlevels)]

    # --- stage dumps (parity tests)
    def blurred_level(self, level, frame=0):
        w, h = self.level_size(level)
        o = np.zeros((h, w), np.uint8)
        _check(self._L.orbx_debug_blurred_level(self._h, frame, level, _p(o), o.strides[0]))
        return o

    def level_candidates(self, level, frame=0):
        n = C.c_int(0)
        _check(self._L.orbx_debug_candidates(self._h, frame, level, None, 0, C.byref(n)))
        o = np.zeros((max(n.value, 1), 3), np.int32)
        _check(self._L.orbx_debug_candidates(self._h, frame, level, _p(o), n.value, C.byref(n)))
        return o[:n.value]

    def level_count(self, level, frame=0):
        n = C.c_int(0)
        _check(self._L.orbx_debug_level_count(self._h, frame, level, C.byref(n)))
        return n.value


class DeviceFrame:
    """orbm_frame (include/orbx.h): undistorted keypoints, descriptors and the 64 x 48 feature grid of a Frame, resident on the device."""

    def __init__(self, matcher, kps=None, desc=None, bounds=None, extractor=None, frame=0, K4=None, dist_coef=None):
        self._L = lib()
        self._m = matcher
        h = C.c_void_p()
        b = [float(v) for v in bounds]
        if extractor is None:
            kps = np.ascontiguousarray(kps, KP_DTYPE)
            desc = _u8(desc).reshape(-1, 32)
            _check(self._L.orbm_frame_create(matcher._h, _p(kps), _p(desc), len(kps), *b, C.byref(h)))
        else:
            K4 = K4 if K4 is not None else (1.0, 1.0, 0.0, 0.0)
            dc = None if dist_coef is None else np.ascontiguousarray(dist_coef, np.float32)
            _check(self._L.orbm_frame_from_extractor(matcher._h, extractor._h, int(frame), *[float(v) for v in K4],
                                                     None if dc is None else _p(dc), 0 if dc is None else len(dc), *b, C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            self._L.orbm_frame_destroy(self._h)
            self._h = None

    __del__ = close

    def __len__(self):
        return self._L.orbm_frame_size(self._h)

    def keypoints(self):
        k = np.zeros(max(len(self), 1), KP_DTYPE)
        _check(self._L.orbm_frame_keypoints(self._h, _p(k)))
        return k[:len(self)]

    def grid(self):
        cs = np.zeros(64 * 48 + 1, np.int32)
        en = np.zeros(max(len(self), 1), np.int32)
        _check(self._L.orbm_frame_grid(self._m._h, self._h, _p(cs), _p(en)))
        return cs, en[:len(self)]


class ExtractorPool:
    """orbx_pool_* (include/orbx.h): one process, every GPU -- a batch of frames sharded over devices in chunks, results in frame order."""

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7, devices=None, chunk=64, lanes=2):
        self._L = lib()
        h = C.c_void_p()
        if devices is None:
            _check(self._L.orbx_pool_create(None, 0, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, chunk, lanes, C.byref(h)))
        else:
            arr = (C.c_int * len(devices))(*devices)
            _check(self._L.orbx_pool_create(arr, len(devices), nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, chunk, lanes, C.byref(h)))
        self._h = h
        self._caps = {}                             # per-frame keypoint capacity of the shapes planned so far

    def close(self):
        if getattr(self, "_h", None):
            self._L.orbx_pool_destroy(self._h)
            self._h = None

    __del__ = close

    def devices(self):
        return self._L.orbx_pool_devices(self._h)

    def set_fma_mode(self, mode):
        _check(self._L.orbx_pool_set_fma_mode(self._h, int(mode)))

    def plan(self, w, h):
        cap = C.c_int(0)
        _check(self._L.orbx_pool_plan(self._h, w, h, C.byref(cap)))
        return cap.value

    def extract_batch(self, images):
        """images: [B, H, W] uint8 array (or a sequence of equally-shaped 2-D arrays) -> list of (kps, desc), frame order."""
        imgs = [_u8(im) for im in images]
        B = len(imgs)
        if B == 0:
            return []
        hh, ww = imgs[0].shape
        cap = self.plan(ww, hh)
        kps = np.zeros((B, cap), KP_DTYPE)
        desc = np.zeros((B, cap, 32), np.uint8)
        n = np.zeros(B, np.int32)
        ptrs = (C.c_void_p * B)(*[im.ctypes.data for im in imgs])
        _check(self._L.orbx_pool_extract_batch(self._h, ptrs, B, ww, hh, imgs[0].strides[0], _p(kps), _p(desc), cap, _p(n)))
        return [(kps[b, :n[b]].copy(), desc[b, :n[b]].copy()) for b in range(B)]

    def submit(self, images):
        """orbx_pool_submit_batch: queue a batch and return a ticket object; several may be in flight.  wait(ticket) -> results."""
        imgs = [_u8(im) for im in images]
        B = len(imgs)
        if B == 0:
            return None
        hh, ww = imgs[0].shape
        cap = self._caps.get((ww, hh))
        if cap is None:
            cap = self._caps[(ww, hh)] = self.plan(ww, hh)          # planning waits for the batches in flight
        kps = np.zeros((B, cap), KP_DTYPE)
        desc = np.zeros((B, cap, 32), np.uint8)
        n = np.zeros(B, np.int32)
        ptrs = (C.c_void_p * B)(*[im.ctypes.data for im in imgs])
        t = C.c_ulonglong(0)
        _check(self._L.orbx_pool_submit_batch(self._h, ptrs, B, ww, hh, imgs[0].strides[0], _p(kps), _p(desc), cap, _p(n), C.byref(t)))
        return (t.value, imgs, kps, desc, n)        # the ticket keeps the frames and the result arrays alive

    def wait(self, ticket):
        if ticket is None:
            return []
        t, _imgs, kps, desc, n = ticket
        _check(self._L.orbx_pool_wait(self._h, t))
        return [(kps[b, :n[b]].copy(), desc[b, :n[b]].copy()) for b in range(len(n))]


class ORBmatcher:
    """ORB_SLAM2::ORBmatcher (include/ORBmatcher.h:37-102): the descriptor-level searches."""

    TH_LOW, TH_HIGH, HISTO_LENGTH = 50, 100, 30        # src/ORBmatcher.cc:37-39

    def __init__(self, nnratio=0.6, checkOri=True, device=-1):
        """device: CUDA ordinal, -1 = the calling thread's current device (contexts are recycled by the library)."""
        self._L = lib()
        h = C.c_void_p()
        _check(self._L.orbm_create(device, C.byref(h)))
        self._h = h
        self.mfNNratio, self.mbCheckOrientation = float(nnratio), bool(checkOri)

    def close(self):
        if getattr(self, "_h", None):
            self._L.orbm_destroy(self._h)
            self._h = None

    __del__ = close

    def stream(self):
        return self._L.orbm_stream(self._h)

    def device(self):
        return self._L.orbm_device(self._h)

    def DescriptorDistance(self, a, b):
        """ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1649-1665); a, b: 32-byte rows or [n, 32] arrays."""
        a, b = _u8(a).reshape(-1, 32), _u8(b).reshape(-1, 32)
        assert a.shape == b.shape
        d = np.zeros(len(a), np.int32)
        _check(self._L.orbm_descriptor_distance(self._h, _p(a), _p(b), len(a), _p(d)))
        return int(d[0]) if len(d) == 1 else d

    def match_bruteforce(self, q, t):
        q, t = _u8(q).reshape(-1, 32), _u8(t).reshape(-1, 32)
        nq = len(q)
        bi, bd, sd = (np.zeros(nq, np.int32) for _ in range(3))
        _check(self._L.orbm_match_bruteforce(self._h, _p(q), nq, _p(t), len(t), _p(bi), _p(bd), _p(sd)))
        return bi, bd, sd

    def match_ratio(self, q, t, th_low=None, nnratio=None):
        q, t = _u8(q).reshape(-1, 32), _u8(t).reshape(-1, 32)
        nq = len(q)
        bi, bd, sd = (np.zeros(nq, np.int32) for _ in range(3))
        acc = np.zeros(nq, np.uint8)
        _check(self._L.orbm_match_ratio(self._h, _p(q), nq, _p(t), len(t), self.TH_LOW if th_low is None else th_low,
                                        self.mfNNratio if nnratio is None else nnratio, _p(bi), _p(bd), _p(sd), _p(acc)))
        return bi, bd, sd, acc.astype(bool)

    def match_bruteforce_device(self, d_q, nq, d_t, nt, d_bi, d_bd, d_sd):
        _check(self._L.orbm_match_bruteforce_device(self._h, C.c_void_p(d_q), nq, C.c_void_p(d_t), nt,
                                                    C.c_void_p(d_bi), C.c_void_p(d_bd), C.c_void_p(d_sd)))

    def match_candidates(self, q, t, off, idx):
        q, t = _u8(q).reshape(-1, 32), _u8(t).reshape(-1, 32)
        off = np.ascontiguousarray(off, np.int32)
        idx = np.ascontiguousarray(idx, np.int32)
        nq = len(q)
        bi, bd, si, sd = (np.zeros(nq, np.int32) for _ in range(4))
        _check(self._L.orbm_match_candidates(self._h, _p(q), nq, _p(t), len(t), _p(off), _p(idx), _p(bi), _p(bd), _p(si), _p(sd)))
        return bi, bd, si, sd

    def candidate_distances(self, q, t, off, idx):
        q, t = _u8(q).reshape(-1, 32), _u8(t).reshape(-1, 32)
        off = np.ascontiguousarray(off, np.int32)
        idx = np.ascontiguousarray(idx, np.int32)
        d = np.zeros(max(len(idx), 1), np.int32)
        _check(self._L.orbm_candidate_distances(self._h, _p(q), len(q), _p(t), len(t), _p(off), _p(idx), _p(d)))
        return d[:len(idx)]

    def grid_candidates(self, kps, bounds, qxy, r, minLevel=-1, maxLevel=-1, q_level=None):
        """Frame::GetFeaturesInArea (src/Frame.cc:342-395) for all queries -> CSR (off, idx).
        r: scalar or per-query radii; q_level: optional per-query level (minLevel = maxLevel = q_level[i])."""
        kps = np.ascontiguousarray(kps, KP_DTYPE)
        qxy = np.ascontiguousarray(qxy, np.float32).reshape(-1, 2)
        nq = len(qxy)
        qr = np.ascontiguousarray(np.broadcast_to(np.asarray(r, np.float32), (nq,)))
        ql = None if q_level is None else np.ascontiguousarray(q_level, np.int32)
        off = np.zeros(nq + 1, np.int32)
        cap = max(64 * nq, 1024)
        for _ in range(2):
            idx = np.zeros(cap, np.int32)
            rc = _check(self._L.orbm_grid_candidates(self._h, _p(kps), len(kps), *[float(b) for b in bounds], _p(qxy), nq, _p(qr),
                                                     minLevel, maxLevel, _p(ql), _p(off), _p(idx), cap), allow=(ERR_CAPACITY,))
            if rc == OK:
                break
            cap = int(off[nq])
        return off, idx[:off[nq]].copy()

    def SearchForInitialization(self, k1, d1, k2, d2, bounds, vbPrevMatched, windowSize=10):
        """ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:405-520).
        Returns (nmatches, vnMatches12, updated vbPrevMatched)."""
        k1 = np.ascontiguousarray(k1, KP_DTYPE)
        k2 = np.ascontiguousarray(k2, KP_DTYPE)
        d1, d2 = _u8(d1).reshape(-1, 32), _u8(d2).reshape(-1, 32)
        prev = np.ascontiguousarray(vbPrevMatched, np.float32).copy()
        m12 = np.zeros(len(k1), np.int32)
        nm = C.c_int(0)
        _check(self._L.orbm_search_for_initialization(self._h, _p(k1), _p(d1), len(k1), _p(k2), _p(d2), len(k2),
                                                      *[float(b) for b in bounds], _p(prev), _p(m12), windowSize,
                                                      self.mfNNratio, int(self.mbCheckOrientation), C.byref(nm)))
        return nm.value, m12, prev

    def SearchByProjection(self, kps, desc, u_right, occupied, bounds, scale_factors, proj_xy, proj_xr, level, view_cos,
                           mp_desc, mp_has_obs, th=1.0):
        """ORBmatcher::SearchByProjection(Frame&, vpMapPoints, th) (src/ORBmatcher.cc:45-129) on plain arrays (see
        include/orbx.h).  Returns (nmatches, feature_match)."""
        kps = np.ascontiguousarray(kps, KP_DTYPE)
        desc, mp_desc = _u8(desc).reshape(-1, 32), _u8(mp_desc).reshape(-1, 32)
        f32 = lambda a: np.ascontiguousarray(a, np.float32)
        u_right, sf, proj_xy, proj_xr, view_cos = f32(u_right), f32(scale_factors), f32(proj_xy).reshape(-1, 2), f32(proj_xr), f32(view_cos)
        occupied, mp_has_obs, level = _u8(occupied), _u8(mp_has_obs), np.ascontiguousarray(level, np.int32)
        fm = np.zeros(max(len(kps), 1), np.int32)
        nm = C.c_int(0)
        _check(self._L.orbm_search_by_projection(self._h, _p(kps), _p(desc), _p(u_right), _p(occupied), len(kps), *[float(b) for b in bounds],
                                                 _p(sf), len(sf), _p(proj_xy), _p(proj_xr), _p(level), _p(view_cos), _p(mp_desc),
                                                 _p(mp_has_obs), len(level), float(th), self.mfNNratio, _p(fm), C.byref(nm)))
        return nm.value, fm[:len(kps)]

    def SearchByProjectionFrame(self, kps, desc, u_right, occupied, bounds, scale_factors, proj_xy, proj_ur, last_octave, last_angle,
                                mp_desc, mp_has_obs, th, mode=0):
        """ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (src/ORBmatcher.cc:1330-1472) on plain arrays.
        mode: 0 neither, 1 forward, 2 backward.  Returns (nmatches, feature_match)."""
        kps = np.ascontiguousarray(kps, KP_DTYPE)
        desc, mp_desc = _u8(desc).reshape(-1, 32), _u8(mp_desc).reshape(-1, 32)
        f32 = lambda a: np.ascontiguousarray(a, np.float32)
        u_right, sf, proj_xy, proj_ur, last_angle = f32(u_right), f32(scale_factors), f32(proj_xy).reshape(-1, 2), f32(proj_ur), f32(last_angle)
        occupied, mp_has_obs, last_octave = _u8(occupied), _u8(mp_has_obs), np.ascontiguousarray(last_octave, np.int32)
        fm = np.zeros(max(len(kps), 1), np.int32)
        nm = C.c_int(0)
        _check(self._L.orbm_search_by_projection_frame(self._h, _p(kps), _p(desc), _p(u_right), _p(occupied), len(kps),
                                                       *[float(b) for b in bounds], _p(sf), len(sf), _p(proj_xy), _p(proj_ur), _p(last_octave),
                                                       _p(last_angle), _p(mp_desc), _p(mp_has_obs), len(last_octave), float(th), int(mode),
                                                       int(self.mbCheckOrientation), _p(fm), C.byref(nm)))
        return nm.value, fm[:len(kps)]

    # ---- the same three searches with the frame side resident on the device (DeviceFrame)
    def SearchForInitializationDev(self, F1, F2, vbPrevMatched, windowSize=10):
        prev = np.ascontiguousarray(vbPrevMatched, np.float32).copy()
        m12 = np.zeros(max(len(F1), 1), np.int32)
        nm = C.c_int(0)
        _check(self._L.orbm_search_for_initialization_dev(self._h, F1._h, F2._h, _p(prev), _p(m12), windowSize, self.mfNNratio,
                                                          int(self.mbCheckOrientation), C.byref(nm)))
        return nm.value, m12[:len(F1)], prev

    def SearchByProjectionDev(self, F, u_right, occupied, scale_factors, proj_xy, proj_xr, level, view_cos, mp_desc, mp_has_obs, th=1.0):
        f32 = lambda a: np.ascontiguousarray(a, np.float32)
        mp_desc = _u8(mp_desc).reshape(-1, 32)
        u_right, sf, proj_xy, proj_xr, view_cos = f32(u_right), f32(scale_factors), f32(proj_xy).reshape(-1, 2), f32(proj_xr), f32(view_cos)
        occupied, mp_has_obs, level = _u8(occupied), _u8(mp_has_obs), np.ascontiguousarray(level, np.int32)
        fm = np.zeros(max(len(F), 1), np.int32)
        nm = C.c_int(0)
        _check(self._L.orbm_search_by_projection_dev(self._h, F._h, _p(u_right), _p(occupied), _p(sf), len(sf), _p(proj_xy), _p(proj_xr), _p(level),
                                                     _p(view_cos), _p(mp_desc), _p(mp_has_obs), len(level), float(th), self.mfNNratio, _p(fm), C.byref(nm)))
        return nm.value, fm[:len(F)]

    def SearchByProjectionFrameDev(self, F, u_right, occupied, scale_factors, proj_xy, proj_ur, last_octave, last_angle, mp_desc, mp_has_obs, th, mode=0):
        f32 = lambda a: np.ascontiguousarray(a, np.float32)
        mp_desc = _u8(mp_desc).reshape(-1, 32)
        u_right, sf, proj_xy, proj_ur, last_angle = f32(u_right), f32(scale_factors), f32(proj_xy).reshape(-1, 2), f32(proj_ur), f32(last_angle)
        occupied, mp_has_obs, last_octave = _u8(occupied), _u8(mp_has_obs), np.ascontiguousarray(last_octave, np.int32)
        fm = np.zeros(max(len(F), 1), np.int32)
        nm = C.c_int(0)
        _check(self._L.orbm_search_by_projection_frame_dev(self._h, F._h, _p(u_right), _p(occupied), _p(sf), len(sf), _p(proj_xy), _p(proj_ur),
                                                           _p(last_octave), _p(last_angle), _p(mp_desc), _p(mp_has_obs), len(last_octave), float(th),
                                                           int(mode), int(self.mbCheckOrientation), _p(fm), C.byref(nm)))
        return nm.value, fm[:len(F)]

    def SearchByBoW(self, k1, d1, valid1, fv1, k2, d2, fv2, valid2=None):
        """ORBmatcher::SearchByBoW (src/ORBmatcher.cc:159-288 when valid2 is None: key frame -> frame; :522-655 otherwise:
        key frame -> key frame).  fv = (node ids ascending, offsets, feature indices) CSR of a DBoW2::FeatureVector.
        Returns (nmatches, matches)."""
        k1, k2 = np.ascontiguousarray(k1, KP_DTYPE), np.ascontiguousarray(k2, KP_DTYPE)
        d1, d2 = _u8(d1).reshape(-1, 32), _u8(d2).reshape(-1, 32)
        i32 = lambda a: np.ascontiguousarray(a, np.int32)
        n1_, o1_, i1_ = [i32(a) for a in fv1]
        n2_, o2_, i2_ = [i32(a) for a in fv2]
        variant = 0 if valid2 is None else 1
        valid1 = _u8(valid1)
        v2 = None if valid2 is None else _u8(valid2)
        out = np.zeros(max(len(k2) if variant == 0 else len(k1), 1), np.int32)
        nm = C.c_int(0)
        _check(self._L.orbm_search_by_bow(self._h, variant, _p(k1), _p(d1), len(k1), _p(valid1), _p(n1_), _p(o1_), _p(i1_), len(n1_),
                                          _p(k2), _p(d2), len(k2), _p(v2), _p(n2_), _p(o2_), _p(i2_), len(n2_),
                                          self.mfNNratio, int(self.mbCheckOrientation), _p(out), C.byref(nm)))
        return nm.value, out[:(len(k2) if variant == 0 else len(k1))]

    def ComputeStereoMatches(self, left, right, kl, dl, kr, dr, bf, maxD, minD=0.0, frame=0):
        """Frame::ComputeStereoMatches (src/Frame.cc:481-655) on the device pyramids of two extractor handles.
        Returns (n surviving matches, mvuRight, mvDepth)."""
        kl = np.ascontiguousarray(kl, KP_DTYPE)
        kr = np.ascontiguousarray(kr, KP_DTYPE)
        dl, dr = _u8(dl).reshape(-1, 32), _u8(dr).reshape(-1, 32)
        uR = np.zeros(len(kl), np.float32)
        dep = np.zeros(len(kl), np.float32)
        nm = C.c_int(0)
        _check(self._L.orbm_stereo_matches(self._h, left._h, right._h, frame, _p(kl), _p(dl), len(kl), _p(kr), _p(dr), len(kr),
                                           bf, minD, maxD, _p(uR), _p(dep), C.byref(nm)))
        return nm.value, uR, dep


def _stereo_batch(self, left, right, bf, maxD, minD=0.0):
    """Frame::ComputeStereoMatches for every pair of the last batch on two extractor handles, device-resident.
    Returns (nmatched[B], uRight[B, cap], depth[B, cap]); entries beyond a pair's left keypoint count are undefined."""
    B = left.last_batch
    cap = max(left.max_keypoints(), 64)
    uR = np.zeros((B, cap), np.float32)
    dep = np.zeros((B, cap), np.float32)
    nm = np.zeros(B, np.int32)
    _check(self._L.orbm_stereo_matches_batch(self._h, left._h, right._h, bf, minD, maxD, _p(uR), _p(dep), cap, _p(nm)))
    return nm, uR, dep


ORBmatcher.ComputeStereoMatchesBatch = _stereo_batch


def _f32(a):
    return np.ascontiguousarray(a, np.float32)


def _i32(a):
    return np.ascontiguousarray(a, np.int32)


def _search_in_area(self, kps, desc, bounds, qxy, qr, qlevel, qdesc, u_right=None, skip=None, qur=None, inv_level_sigma2=None, nlevels=8):
    """Window + level + (optional chi-square) + Hamming argmin for a batch of projected points, one kernel
    (the loop shared by Fuse and SearchBySim3, src/ORBmatcher.cc:886-935).  Returns (best_idx, best_dist)."""
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    desc, qdesc = _u8(desc).reshape(-1, 32), _u8(qdesc).reshape(-1, 32)
    qxy, qr, qlevel = _f32(qxy), _f32(qr), _i32(qlevel)
    ur = None if u_right is None else _f32(u_right)
    sk = None if skip is None else _u8(skip)
    qu = None if qur is None else _f32(qur)
    inv = None if inv_level_sigma2 is None else _f32(inv_level_sigma2)
    nq = len(qlevel)
    bi, bd = np.zeros(max(nq, 1), np.int32), np.zeros(max(nq, 1), np.int32)
    _check(self._L.orbm_search_in_area(self._h, _p(kps), _p(desc), _p(ur), _p(sk), len(kps), *[float(b) for b in bounds], _p(qxy), _p(qu),
                                       _p(qr), _p(qlevel), _p(qdesc), nq, _p(inv), nlevels if inv is None else len(inv), _p(bi), _p(bd)))
    return bi[:nq], bd[:nq]


def _fuse(self, kps, desc, u_right, bounds, sf, inv_level_sigma2, proj_xy, proj_ur, level, mp_desc, th=3.0, stereo_chi2=True):
    """ORBmatcher::Fuse(pKF, vpMapPoints, th) (src/ORBmatcher.cc:827-975) or, with stereo_chi2=False, Fuse(pKF, Scw, vpPoints, th,
    vpReplacePoint) (:977-1102).  Returns (nFused, best_idx[point] or -1)."""
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    desc, mp_desc = _u8(desc).reshape(-1, 32), _u8(mp_desc).reshape(-1, 32)
    u_right, sf, inv, proj_xy, proj_ur, level = _f32(u_right), _f32(sf), _f32(inv_level_sigma2), _f32(proj_xy), _f32(proj_ur), _i32(level)
    out = np.zeros(max(len(level), 1), np.int32)
    nf = C.c_int(0)
    _check(self._L.orbm_fuse(self._h, _p(kps), _p(desc), _p(u_right), len(kps), *[float(b) for b in bounds], _p(sf), _p(inv), len(sf),
                             _p(proj_xy), _p(proj_ur), _p(level), _p(mp_desc), len(level), float(th), int(stereo_chi2), _p(out), C.byref(nf)))
    return nf.value, out[:len(level)]


def _search_by_sim3(self, k1, d1, k2, d2, bounds, sf, proj12_xy, level12, mp1_desc, proj21_xy, level21, mp2_desc, th=7.5):
    """ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1104-1328).  Returns (nFound, match12)."""
    k1, k2 = np.ascontiguousarray(k1, KP_DTYPE), np.ascontiguousarray(k2, KP_DTYPE)
    d1, d2, m1, m2 = _u8(d1).reshape(-1, 32), _u8(d2).reshape(-1, 32), _u8(mp1_desc).reshape(-1, 32), _u8(mp2_desc).reshape(-1, 32)
    sf, p12, p21, l12, l21 = _f32(sf), _f32(proj12_xy), _f32(proj21_xy), _i32(level12), _i32(level21)
    out = np.zeros(max(len(k1), 1), np.int32)
    nf = C.c_int(0)
    _check(self._L.orbm_search_by_sim3(self._h, _p(k1), _p(d1), len(k1), _p(k2), _p(d2), len(k2), *[float(b) for b in bounds], _p(sf), len(sf),
                                       _p(p12), _p(l12), _p(m1), _p(p21), _p(l21), _p(m2), float(th), _p(out), C.byref(nf)))
    return nf.value, out[:len(k1)]


def _search_by_projection_sim3(self, kps, desc, matched, bounds, sf, proj_xy, level, mp_desc, th=10.0):
    """ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (src/ORBmatcher.cc:290-403)."""
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    desc, mp_desc, matched = _u8(desc).reshape(-1, 32), _u8(mp_desc).reshape(-1, 32), _u8(matched)
    sf, proj_xy, level = _f32(sf), _f32(proj_xy), _i32(level)
    fm = np.zeros(max(len(kps), 1), np.int32)
    nm = C.c_int(0)
    _check(self._L.orbm_search_by_projection_sim3(self._h, _p(kps), _p(desc), _p(matched), len(kps), *[float(b) for b in bounds], _p(sf), len(sf),
                                                  _p(proj_xy), _p(level), _p(mp_desc), len(level), float(th), _p(fm), C.byref(nm)))
    return nm.value, fm[:len(kps)]


def _search_by_projection_kf(self, kps, desc, occupied, bounds, sf, proj_xy, level, kf_angle, mp_desc, th=10.0, orb_dist=100):
    """ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:1474-1601)."""
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    desc, mp_desc, occupied = _u8(desc).reshape(-1, 32), _u8(mp_desc).reshape(-1, 32), _u8(occupied)
    sf, proj_xy, level, kf_angle = _f32(sf), _f32(proj_xy), _i32(level), _f32(kf_angle)
    fm = np.zeros(max(len(kps), 1), np.int32)
    nm = C.c_int(0)
    _check(self._L.orbm_search_by_projection_kf(self._h, _p(kps), _p(desc), _p(occupied), len(kps), *[float(b) for b in bounds], _p(sf), len(sf),
                                                _p(proj_xy), _p(level), _p(kf_angle), _p(mp_desc), len(level), float(th), int(orb_dist),
                                                int(self.mbCheckOrientation), _p(fm), C.byref(nm)))
    return nm.value, fm[:len(kps)]


def _search_for_triangulation(self, k1, d1, has_mp1, u_right1, fv1, k2, d2, has_mp2, u_right2, fv2, F12, epipole, sf2, sigma2_2,
                              only_stereo=False):
    """ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:657-825).  Returns (nmatches, matches12)."""
    k1, k2 = np.ascontiguousarray(k1, KP_DTYPE), np.ascontiguousarray(k2, KP_DTYPE)
    d1, d2, h1, h2 = _u8(d1).reshape(-1, 32), _u8(d2).reshape(-1, 32), _u8(has_mp1), _u8(has_mp2)
    u1, u2, F, sf2, sg = _f32(u_right1), _f32(u_right2), _f32(np.asarray(F12).reshape(9)), _f32(sf2), _f32(sigma2_2)
    n1_, o1_, i1_ = [_i32(a) for a in fv1]
    n2_, o2_, i2_ = [_i32(a) for a in fv2]
    out = np.zeros(max(len(k1), 1), np.int32)
    nm = C.c_int(0)
    _check(self._L.orbm_search_for_triangulation(self._h, _p(k1), _p(d1), len(k1), _p(h1), _p(u1), _p(n1_), _p(o1_), _p(i1_), len(n1_),
                                                 _p(k2), _p(d2), len(k2), _p(h2), _p(u2), _p(n2_), _p(o2_), _p(i2_), len(n2_),
                                                 _p(F), float(epipole[0]), float(epipole[1]), _p(sf2), _p(sg), len(sf2),
                                                 int(only_stereo), int(self.mbCheckOrientation), _p(out), C.byref(nm)))
    return nm.value, out[:len(k1)]


def _distinctive_descriptors(self, desc, off):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:252-312) for many MapPoints: the observations of point p are
    rows off[p]..off[p+1] of desc.  Returns, per point, the index within its rows of the most distinctive descriptor (-1: none)."""
    desc, off = _u8(desc).reshape(-1, 32), _i32(off)
    npts = len(off) - 1
    best = np.zeros(max(npts, 1), np.int32)
    _check(self._L.orbm_distinctive_descriptors(self._h, _p(desc), _p(off), npts, _p(best)))
    return best[:npts]


ORBmatcher.ComputeDistinctiveDescriptors = _distinctive_descriptors
def _undistort_points(self, pts, K4, dist):
    """cv::undistortPoints(pts, K, dist, noArray(), K) as used by Frame (src/Frame.cc:438, :464); K4 = (fx, fy, cx, cy)."""
    pts, dist = _f32(pts).reshape(-1, 2), _f32(dist).reshape(-1)
    out = np.zeros_like(pts)
    _check(self._L.orbm_undistort_points(self._h, _p(pts), len(pts), *[float(np.float32(v)) for v in K4], _p(dist), len(dist), _p(out)))
    return out


def _undistort_keypoints(self, kps, K4, dist):
    """Frame::UndistortKeyPoints (src/Frame.cc:419-449) -> mvKeysUn."""
    kps, dist = np.ascontiguousarray(kps, KP_DTYPE), _f32(dist).reshape(-1)
    out = np.zeros_like(kps)
    _check(self._L.orbm_undistort_keypoints(self._h, _p(kps), len(kps), *[float(np.float32(v)) for v in K4], _p(dist), len(dist), _p(out)))
    return out


def _image_bounds(self, cols, rows, K4, dist):
    """Frame::ComputeImageBounds (src/Frame.cc:451-479) -> (mnMinX, mnMaxX, mnMinY, mnMaxY)."""
    dist = _f32(dist).reshape(-1)
    out = np.zeros(4, np.float32)
    _check(self._L.orbm_image_bounds(self._h, cols, rows, *[float(np.float32(v)) for v in K4], _p(dist), len(dist), _p(out)))
    return out


def _stereo_from_rgbd(self, kps, kps_un, im_depth, bf):
    """Frame::ComputeStereoFromRGBD (src/Frame.cc:658-679) -> (mvuRight, mvDepth)."""
    kps, kps_un, im = np.ascontiguousarray(kps, KP_DTYPE), np.ascontiguousarray(kps_un, KP_DTYPE), _f32(im_depth)
    uR, dep = np.zeros(len(kps), np.float32), np.zeros(len(kps), np.float32)
    _check(self._L.orbm_stereo_from_rgbd(self._h, _p(kps), _p(kps_un), len(kps), _p(im), im.shape[1], im.shape[0], im.strides[0], float(bf),
                                         _p(uR), _p(dep)))
    return uR, dep


ORBmatcher.UndistortPoints = _undistort_points
ORBmatcher.UndistortKeyPoints = _undistort_keypoints
ORBmatcher.ComputeImageBounds = _image_bounds
ORBmatcher.ComputeStereoFromRGBD = _stereo_from_rgbd
ORBmatcher.SearchInArea = _search_in_area
ORBmatcher.last_launches = lambda self: self._L.orbm_last_launches(self._h)
ORBmatcher.Fuse = _fuse
ORBmatcher.SearchBySim3 = _search_by_sim3
ORBmatcher.SearchByProjectionSim3 = _search_by_projection_sim3
ORBmatcher.SearchByProjectionKF = _search_by_projection_kf
ORBmatcher.SearchForTriangulation = _search_for_triangulation


class ORBVocabulary:
    """Mirror of the reference's ORBVocabulary (include/ORBVocabulary.h:31-32, DBoW2::TemplatedVocabulary<FORB>) over the orbv_* C ABI:
    the tree lives on the device, transform() descends all descriptors of a frame in one launch."""
    L1_NORM, L2_NORM, CHI_SQUARE, KL, BHATTACHARYYA, DOT_PRODUCT = range(6)
    TF_IDF, TF, IDF, BINARY = range(4)

    def __init__(self, k, L, parent, desc, weight, scoring=0, weighting=0, device=0):
        self._L = lib()
        self._h = C.c_void_p()
        self.k, self.L, self.scoring, self.weighting = k, L, scoring, weighting
        parent, desc, weight = _i32(parent), _u8(desc).reshape(-1, 32), np.ascontiguousarray(weight, np.float64)
        _check(self._L.orbv_create(device, k, L, scoring, weighting, _p(parent), _p(desc), _p(weight), len(parent), C.byref(self._h)))

    def __del__(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self._L.orbv_destroy(self._h)
            self._h = C.c_void_p()

    @staticmethod
    def parse_text(filename):
        """The ORBvoc.txt format of loadFromTextFile (TemplatedVocabulary.h:1351-1436): header `k L scoring weighting`, then one line
        per node: parent, is-leaf flag, 32 descriptor bytes, weight.  Returns (k, L, scoring, weighting, parent, desc, weight)."""
        with open(filename) as f:
            k, L, scoring, weighting = [int(t) for t in f.readline().split()[:4]]
            parent, desc, weight = [0], [np.zeros(32, np.uint8)], [0.0]
            for line in f:
                t = line.split()
                if len(t) < 35:
                    continue
                parent.append(int(t[0]))
                desc.append(np.array([int(b) for b in t[2:34]], np.uint8))
                weight.append(float(t[34]))
        return k, L, scoring, weighting, np.asarray(parent, np.int32), np.stack(desc), np.asarray(weight, np.float64)

    @classmethod
    def loadFromTextFile(cls, filename, device=0):
        k, L, scoring, weighting, parent, desc, weight = cls.parse_text(filename)
        return cls(k, L, parent, desc, weight, scoring, weighting, device)

    def size(self):
        return self._L.orbv_size(self._h)

    def transform_raw(self, desc, levelsup=4):
        """Per-descriptor (word id, weight, node id at level L - levelsup)."""
        desc = _u8(desc).reshape(-1, 32)
        n = len(desc)
        wid, nid, w = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.float64)
        _check(self._L.orbv_transform_raw(self._h, _p(desc), n, levelsup, _p(wid), _p(w), _p(nid)))
        return wid[:n], w[:n], nid[:n]

    def transform(self, desc, levelsup=4):
        """transform(features, BowVec, FeatVec, levelsup) (src/Frame.cc:410-417) -> ((bow ids, bow values), (fv nodes, offsets, indices));
        the second tuple is the CSR form ORBmatcher.SearchByBoW takes."""
        desc = _u8(desc).reshape(-1, 32)
        n = len(desc)
        bid, bval = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.float64)
        fn, fo, fi = np.zeros(max(n, 1), np.int32), np.zeros(n + 1, np.int32), np.zeros(max(n, 1), np.int32)
        nb, nf = C.c_int(0), C.c_int(0)
        _check(self._L.orbv_transform(self._h, _p(desc), n, levelsup, _p(bid), _p(bval), C.byref(nb), _p(fn), _p(fo), _p(fi), C.byref(nf)))
        return (bid[:nb.value], bval[:nb.value]), (fn[:nf.value], fo[:nf.value + 1], fi[:fo[nf.value]])


def version():
    return lib().orbx_version().decode()


class PinnedArray:
    """A page-locked numpy array from orbx_alloc_pinned (include/orbx.h); write_combined=True for buffers the host only
    writes (frames on their way to the device).  Keep the object alive while `.array` is in use."""

    def __init__(self, shape, dtype=np.uint8, write_combined=False):
        self.nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = C.c_void_p()
        _check(lib().orbx_alloc_pinned(C.c_size_t(max(self.nbytes, 1)), int(write_combined), C.byref(p)))
        self.ptr = p.value
        buf = (C.c_uint8 * max(self.nbytes, 1)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def free(self):
        if self.ptr:
            self.array = None
            lib().orbx_free_pinned(C.c_void_p(self.ptr))
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass
