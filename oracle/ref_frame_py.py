"""ctypes binding of oracle/_ref/libframe_ref.so: the UNMODIFIED reference Frame (/root/reference/src/Frame.cc with src/ORBmatcher.cc,
src/ORBextractor.cc and the vendored Thirdparty/DBoW2, compiled where they lie over oracle/ref_shim/; entry points:
oracle/ref_shim/frame_ref_api.cpp).

TEST INFRASTRUCTURE, NOT PRODUCT: only tests/ and the cpu_baseline / --impl reference legs of bench.py (stereo configurations) may import
this module.  The library is built by `make -C oracle _ref` in the container that
holds /root/reference; the GPU box receives the built file with the snapshot.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("ORB_REFERENCE_ROOT", "/root/reference")
KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
PATH = os.path.join(_HERE, "_ref", "libframe_ref.so")
# The same C interface over the reference's Frame.cc compiled against THIS REPOSITORY'S ORBextractor.h / ORBmatcher.h (oracle/Makefile, _ref_tests):
# liborbx.so behind it (GPU), or the C ABI served by the CPU oracle.
DROPIN_PATH = os.path.join(_HERE, "_ref", "libframe_dropin.so")
DROPIN_CPU_PATH = os.path.join(_HERE, "_ref", "libframe_dropin_cpu.so")
_libs = {}


def available():
    if not os.path.exists(PATH) and os.path.exists(os.path.join(REF_ROOT, "src", "Frame.cc")):
        subprocess.check_call(["make", "-C", _HERE, "liborb_oracle.so", "_ref/libframe_ref.so", "REF=" + REF_ROOT])
    return os.path.exists(PATH)


def dropin_available(cpu):
    path = DROPIN_CPU_PATH if cpu else DROPIN_PATH
    if os.path.exists(os.path.join(REF_ROOT, "src", "Frame.cc")) and (cpu or os.path.exists(os.path.join(_HERE, "..", "orb_slam2v2-1_b200", "orbx", "liborbx.so"))):
        subprocess.check_call(["make", "-C", _HERE, "liborb_oracle.so", "_ref/" + os.path.basename(path), "REF=" + REF_ROOT], stdout=subprocess.DEVNULL)
    return os.path.exists(path)


def lib(path=None):
    path = path or PATH
    if path not in _libs:
        if path == PATH and not available():
            raise RuntimeError("oracle/_ref/libframe_ref.so is missing and /root/reference is not here to build it")
        L = C.CDLL(path)
        vp, i, f, d = C.c_void_p, C.c_int, C.c_float, C.c_double
        L.reff_frame_create.restype = vp
        L.reff_frame_create.argtypes = [i, vp, vp, vp, i, i, i, f, i, i, i, vp, vp, i, f, f, vp]
        L.reff_frame_destroy.argtypes = [vp]
        for name in ("reff_frame_n", "reff_frame_n_right"):
            getattr(L, name).restype = i
            getattr(L, name).argtypes = [vp]
        L.reff_frame_mb.restype = f
        L.reff_frame_mb.argtypes = [vp]
        L.reff_frame_get.argtypes = [vp] * 8
        L.reff_frame_get_right.argtypes = [vp] * 3
        L.reff_frame_scales.restype = i
        L.reff_frame_scales.argtypes = [vp] * 5
        L.reff_frame_level.restype = i
        L.reff_frame_level.argtypes = [vp, i, i, vp, vp, vp, vp]
        L.reff_frame_features_in_area.restype = i
        L.reff_frame_features_in_area.argtypes = [vp, f, f, f, i, i, vp, i]
        L.reff_frame_grid_cell.restype = i
        L.reff_frame_grid_cell.argtypes = [vp, i, i, vp, i]
        L.reff_frame_pos_in_grid.restype = i
        L.reff_frame_pos_in_grid.argtypes = [vp, f, f, vp, vp]
        if hasattr(L, "reff_frame_is_in_frustum"):
            L.reff_frame_is_in_frustum.argtypes = [vp, vp, vp, vp, vp, vp, i, f, vp]
        L.reff_frame_unproject_stereo.restype = i
        L.reff_frame_unproject_stereo.argtypes = [vp, i, vp]
        L.reff_voc_load_text.restype = vp
        L.reff_voc_load_text.argtypes = [C.c_char_p]
        L.reff_voc_destroy.argtypes = [vp]
        L.reff_voc_size.restype = i
        L.reff_voc_size.argtypes = [vp]
        L.reff_frame_compute_bow.restype = i
        L.reff_frame_compute_bow.argtypes = [vp] * 7
        L.reff_voc_transform.restype = i
        L.reff_voc_transform.argtypes = [vp, vp, i, i, vp, vp, vp, vp, vp, vp]
        L.reff_voc_score.restype = d
        L.reff_voc_score.argtypes = [vp, vp, vp, i, vp, vp, i]
        L.reff_voc_load_binary.restype = vp
        L.reff_voc_load_binary.argtypes = [C.c_char_p]
        if hasattr(L, "reff_voc_save_binary"):
            L.reff_voc_save_binary.argtypes = [vp, C.c_char_p]
        L.reff_stereo_create.restype = vp
        L.reff_stereo_create.argtypes = [i, f, i, i, i, vp, f, f]
        L.reff_stereo_destroy.argtypes = [vp]
        L.reff_stereo_run.restype = i
        L.reff_stereo_run.argtypes = [vp, vp, vp, i, i, vp]
        _libs[path] = L
    return _libs[path]


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class RefVocabulary:
    """ORBVocabulary of the reference (DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>) loaded with its own loadFromTextFile."""

    def __init__(self, path, libpath=None, binary=False):
        self.L = lib(libpath)
        self.h = (self.L.reff_voc_load_binary if binary else self.L.reff_voc_load_text)(os.fsencode(path))
        if not self.h:
            raise RuntimeError("loadFrom%sFile failed: %s" % ("Binary" if binary else "Text", path))

    def save_binary(self, path):
        """DBoW2's saveToBinaryFile (the reference build only)."""
        self.L.reff_voc_save_binary(self.h, os.fsencode(path))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.reff_voc_destroy(self.h)
            self.h = None

    def size(self):
        return self.L.reff_voc_size(self.h)

    def transform(self, desc, levelsup=4):
        """-> (bow ids, bow values), (fv nodes, fv offsets, fv feature indices), both in map order"""
        desc = np.ascontiguousarray(desc, np.uint8)
        n = len(desc)
        bid, bval = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.float64)
        fn, fo, fi = np.zeros(max(n, 1), np.int32), np.zeros(n + 1, np.int32), np.zeros(max(n, 1), np.int32)
        nfv = C.c_int(0)
        nb = self.L.reff_voc_transform(self.h, _p(desc), n, levelsup, _p(bid), _p(bval), _p(fn), _p(fo), _p(fi), C.byref(nfv))
        return (bid[:nb], bval[:nb]), (fn[:nfv.value], fo[:nfv.value + 1], fi[:fo[nfv.value]])

    def score(self, a, b):
        ia, va = np.ascontiguousarray(a[0], np.int32), np.ascontiguousarray(a[1], np.float64)
        ib, vb = np.ascontiguousarray(b[0], np.int32), np.ascontiguousarray(b[1], np.float64)
        return self.L.reff_voc_score(self.h, _p(ia), _p(va), len(ia), _p(ib), _p(vb), len(ib))


class RefStereoPipeline:
    """Two reference extractors kept across pairs; run(left, right) builds the reference Frame of one rectified pair (extraction of both
    images, grid, ComputeStereoMatches) and returns (N, features with a stereo match).  The timed CPU baseline of the stereo configurations."""

    def __init__(self, K4, bf, thDepth=35.0, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniTh=20, minTh=7, libpath=None):
        self.L = lib(libpath)
        K4 = np.ascontiguousarray(K4, np.float32)
        self.h = self.L.reff_stereo_create(nfeatures, scaleFactor, nlevels, iniTh, minTh, _p(K4), bf, thDepth)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.reff_stereo_destroy(self.h)
            self.h = None

    def run(self, left, right):
        left, right = np.ascontiguousarray(left, np.uint8), np.ascontiguousarray(right, np.uint8)
        counts = np.zeros(2, np.int32)
        if self.L.reff_stereo_run(self.h, _p(left), _p(right), left.shape[1], left.shape[0], _p(counts)) != 0:
            raise RuntimeError("the reference Frame threw")
        return int(counts[0]), int(counts[1])


class RefFrame:
    """A reference Frame built by the reference's constructors.  kind: "mono" (Frame.cc:174), "stereo" (mono constructor for the left image +
    ExtractORB(1, right) + ComputeStereoMatches(), see frame_ref_api.cpp on why), "rgbd" (Frame.cc:119)."""

    def __init__(self, kind, im, K4, dist, bf, thDepth=35.0, im_right=None, depth=None, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniTh=20, minTh=7,
                 voc=None, libpath=None):
        self.L = lib(libpath)
        self.im = np.ascontiguousarray(im, np.uint8)
        self.im_right = None if im_right is None else np.ascontiguousarray(im_right, np.uint8)
        self.depth = None if depth is None else np.ascontiguousarray(depth, np.float32)
        self.voc = voc
        h, w = self.im.shape
        K4, dist = np.ascontiguousarray(K4, np.float32), np.ascontiguousarray(dist, np.float32)
        k = {"mono": 0, "stereo": 1, "rgbd": 2}[kind]
        self.h = self.L.reff_frame_create(k, _p(self.im), None if self.im_right is None else _p(self.im_right),
                                          None if self.depth is None else _p(self.depth), w, h, nfeatures, scaleFactor, nlevels, iniTh, minTh,
                                          _p(K4), _p(dist), len(dist), bf, thDepth, voc.h if voc is not None else None)
        if not self.h:
            raise RuntimeError("the reference Frame constructor threw")
        self.N = self.L.reff_frame_n(self.h)
        n = max(self.N, 1)
        self.keys, self.keys_un = np.zeros(n, KP_DTYPE), np.zeros(n, KP_DTYPE)
        self.desc = np.zeros((n, 32), np.uint8)
        self.u_right, self.depth_out = np.zeros(n, np.float32), np.zeros(n, np.float32)
        self.bounds, self.grid_inv = np.zeros(4, np.float32), np.zeros(2, np.float32)
        self.L.reff_frame_get(self.h, _p(self.keys), _p(self.keys_un), _p(self.desc), _p(self.u_right), _p(self.depth_out), _p(self.bounds), _p(self.grid_inv))
        for name in ("keys", "keys_un", "desc", "u_right", "depth_out"):
            setattr(self, name, getattr(self, name)[:self.N])
        self.mb = self.L.reff_frame_mb(self.h)
        nr = self.L.reff_frame_n_right(self.h)
        self.keys_right, self.desc_right = np.zeros(max(nr, 1), KP_DTYPE), np.zeros((max(nr, 1), 32), np.uint8)
        if nr:
            self.L.reff_frame_get_right(self.h, _p(self.keys_right), _p(self.desc_right))
        self.keys_right, self.desc_right = self.keys_right[:nr], self.desc_right[:nr]
        self.sf, self.isf, self.sigma2, self.isigma2 = (np.zeros(nlevels, np.float32) for _ in range(4))
        self.L.reff_frame_scales(self.h, _p(self.sf), _p(self.isf), _p(self.sigma2), _p(self.isigma2))
        self.nlevels = nlevels

    def __del__(self):
        if getattr(self, "h", None):
            self.L.reff_frame_destroy(self.h)
            self.h = None

    def pyramid(self, side=0):
        """mvImagePyramid of the left / right extractor as numpy views (valid while the frame lives)."""
        out = []
        for l in range(self.nlevels):
            w, h, ptr, step = C.c_int(), C.c_int(), C.c_void_p(), C.c_size_t()
            if self.L.reff_frame_level(self.h, side, l, C.byref(w), C.byref(h), C.byref(ptr), C.byref(step)) != 0:
                raise RuntimeError("no such pyramid level")
            buf = (C.c_uint8 * (step.value * h.value)).from_address(ptr.value)
            out.append(np.lib.stride_tricks.as_strided(np.frombuffer(buf, np.uint8), (h.value, w.value), (step.value, 1)))
        return out

    def features_in_area(self, x, y, r, minLevel=-1, maxLevel=-1):
        buf = np.zeros(max(self.N, 1), np.int32)
        n = self.L.reff_frame_features_in_area(self.h, x, y, r, minLevel, maxLevel, _p(buf), len(buf))
        return buf[:n].copy()

    def grid_cell(self, col, row):
        buf = np.zeros(max(self.N, 1), np.int32)
        n = self.L.reff_frame_grid_cell(self.h, col, row, _p(buf), len(buf))
        return buf[:n].copy()

    def pos_in_grid(self, x, y):
        px, py = C.c_int(), C.c_int()
        ok = self.L.reff_frame_pos_in_grid(self.h, x, y, C.byref(px), C.byref(py))
        return bool(ok), px.value, py.value

    def is_in_frustum(self, Tcw, pos, normal, min_dist, max_dist, cos_limit=0.5):
        Tcw = np.ascontiguousarray(Tcw, np.float32).reshape(16)
        pos, normal = np.ascontiguousarray(pos, np.float32).reshape(-1, 3), np.ascontiguousarray(normal, np.float32).reshape(-1, 3)
        mn, mx = np.ascontiguousarray(min_dist, np.float32), np.ascontiguousarray(max_dist, np.float32)
        out = np.zeros((len(pos), 6), np.float32)
        self.L.reff_frame_is_in_frustum(self.h, _p(Tcw), _p(pos), _p(normal), _p(mn), _p(mx), len(pos), cos_limit, _p(out))
        return out

    def unproject_stereo(self, i):
        xyz = np.zeros(3, np.float32)
        return xyz if self.L.reff_frame_unproject_stereo(self.h, i, _p(xyz)) else None

    def compute_bow(self):
        n = max(self.N, 1)
        bid, bval = np.zeros(n, np.int32), np.zeros(n, np.float64)
        fn, fo, fi = np.zeros(n, np.int32), np.zeros(n + 1, np.int32), np.zeros(n, np.int32)
        nfv = C.c_int(0)
        nb = self.L.reff_frame_compute_bow(self.h, _p(bid), _p(bval), _p(fn), _p(fo), _p(fi), C.byref(nfv))
        if nb < 0:
            raise RuntimeError("frame without a vocabulary")
        return (bid[:nb], bval[:nb]), (fn[:nfv.value], fo[:nfv.value + 1], fi[:fo[nfv.value]])
