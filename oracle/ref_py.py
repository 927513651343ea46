"""ctypes binding of oracle/_ref/liborb_ref*.so: the UNMODIFIED reference extractor
(/root/reference/src/ORBextractor.cc compiled where it lies over oracle/ref_shim/).

TEST INFRASTRUCTURE, NOT PRODUCT: only tests/, __graft_entry__ and the cpu_baseline / --impl reference
legs of bench.py may import this module.  The libraries are built by `make -C oracle _ref` in the
container that holds /root/reference; the GPU box receives the built files with the snapshot.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("ORB_REFERENCE_ROOT", "/root/reference")
KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])

_PATHS = {"fma": os.path.join(_HERE, "_ref", "liborb_ref.so"),          # the reference's own flags (contracted :119-120)
          "nofma": os.path.join(_HERE, "_ref", "liborb_ref_nofma.so")}  # same source, -ffp-contract=off
_libs = {}


def can_build():
    return os.path.exists(os.path.join(REF_ROOT, "src", "ORBextractor.cc"))


def build(force=False):
    """Compile the reference where it lies (needs /root/reference). Returns True if both libraries exist afterwards."""
    if can_build() and (force or not all(os.path.exists(p) for p in _PATHS.values())):
        subprocess.check_call(["make", "-C", _HERE, "_ref", "REF=" + REF_ROOT] + (["-B"] if force else []))
    return available()


def available():
    return all(os.path.exists(p) for p in _PATHS.values())


def lib(variant="fma"):
    if variant not in _libs:
        if not os.path.exists(_PATHS[variant]):
            build()
        L = C.CDLL(_PATHS[variant])
        vp, i, f, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
        L.orbr_build_info.restype = C.c_char_p
        L.orbr_set_allocator.argtypes = [i]
        L.orbr_bump_allocations.restype = C.c_long
        L.orbr_create.restype = vp
        L.orbr_create.argtypes = [i, f, i, i, i]
        L.orbr_destroy.argtypes = [vp]
        L.orbr_extract.restype = i
        L.orbr_extract.argtypes = [vp, vp, i, i, sz, vp, vp, i]
        L.orbr_levels.restype = i
        L.orbr_levels.argtypes = [vp]
        L.orbr_scale_tables.argtypes = [vp, vp, vp, vp, vp]
        L.orbr_accessor_tables.argtypes = [vp, vp, vp, vp, vp, vp, vp]
        L.orbr_features_per_level.argtypes = [vp, vp]
        L.orbr_umax.argtypes = [vp, vp]
        L.orbr_level_size.argtypes = [vp, i, vp, vp]
        L.orbr_level_padded.argtypes = [vp, i, vp]
        L.orbr_distribute.restype = i
        L.orbr_distribute.argtypes = [vp, vp, i, i, i, i, i, i, vp, i]
        _libs[variant] = L
    return _libs[variant]


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class RefExtractor:
    """ORB_SLAM2::ORBextractor itself (include/ORBextractor.h:45-111 of the reference).

    allocator: 0 = malloc/free as a stock build (heap-address tie-break at ORBextractor.cc:684 follows the heap's history),
               1 = monotone arena (addresses grow in allocation order; see oracle/ref_shim/ref_api.cpp).
    variant:  "fma" = the reference's flags, "nofma" = the same with -ffp-contract=off.
    """

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7, allocator=1, variant="fma"):
        self.L = lib(variant)
        self.nlevels = nlevels
        self.nfeatures = nfeatures
        self.allocator = allocator
        self.h = self.L.orbr_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orbr_destroy(self.h)
            self.h = None

    def __call__(self, img):
        img = np.ascontiguousarray(img, dtype=np.uint8)
        hh, ww = img.shape
        cap = max(4 * self.nfeatures + 64 * self.nlevels, 1024)
        kps = np.zeros(cap, dtype=KP_DTYPE)
        desc = np.zeros((cap, 32), dtype=np.uint8)
        self.L.orbr_set_allocator(self.allocator)
        n = self.L.orbr_extract(self.h, _p(img), ww, hh, img.strides[0], _p(kps), _p(desc), cap)
        if n < 0:
            raise RuntimeError("the reference extractor threw")
        assert n <= cap
        return kps[:n].copy(), desc[:n].copy()

    def scale_tables(self):
        out = [np.zeros(self.nlevels, dtype=np.float32) for _ in range(4)]
        self.L.orbr_scale_tables(self.h, *[_p(o) for o in out])
        return out

    def accessor_tables(self):
        out = [np.zeros(self.nlevels, dtype=np.float32) for _ in range(4)]
        lv, sf = C.c_int(), C.c_float()
        self.L.orbr_accessor_tables(self.h, *[_p(o) for o in out], C.byref(lv), C.byref(sf))
        return out, lv.value, sf.value

    def features_per_level(self):
        o = np.zeros(self.nlevels, dtype=np.int32)
        self.L.orbr_features_per_level(self.h, _p(o))
        return o

    def umax(self):
        o = np.zeros(16, dtype=np.int32)
        self.L.orbr_umax(self.h, _p(o))
        return o

    def level_size(self, l):
        w, h = C.c_int(), C.c_int()
        self.L.orbr_level_size(self.h, l, C.byref(w), C.byref(h))
        return w.value, h.value

    def level_padded(self, l):
        w, h = self.level_size(l)
        o = np.zeros((h + 38, w + 38), dtype=np.uint8)
        self.L.orbr_level_padded(self.h, l, _p(o))
        return o

    def distribute(self, xyr, minX, maxX, minY, maxY, N):
        """ORBextractor::DistributeOctTree alone: indices of the kept candidates, in output order."""
        xyr = np.ascontiguousarray(xyr, dtype=np.int32)
        n = len(xyr)
        out = np.zeros(max(n, 1), dtype=np.int32)
        self.L.orbr_set_allocator(self.allocator)
        m = self.L.orbr_distribute(self.h, _p(xyr), n, minX, maxX, minY, maxY, N, _p(out), n)
        if m < 0:
            raise RuntimeError("the reference DistributeOctTree threw")
        return out[:m].copy()
