"""Second, independent oracle: the reference's ORBextractor orchestration re-stated in Python
on top of the REAL OpenCV primitives of the cv2 wheel (4.13.0 here).

TEST INFRASTRUCTURE, NOT PRODUCT.  It exists to pin oracle/liborb_oracle.so: both must agree
byte for byte on full frames (tests/test_oracle_vs_cv2.py).  Follows reference
src/ORBextractor.cc: ctor :410-470, ComputePyramid :1107-1132, ComputeKeyPointsOctTree :765-853,
DistributeOctTree :539-763, IC_Angle :77-104, computeOrbDescriptor :108-147, operator() :1043-1105.
The tie-break of the size sort (:684) is creation order, later first (see orb_octree.cpp).
"""
import math
import os
import numpy as np
import cv2

EDGE = 19
HALF_PATCH = 15
f32 = np.float32


def _pattern():
    here = os.path.dirname(os.path.abspath(__file__))
    txt = open(os.path.join(here, "orb_pattern.inc")).read().split("\n", 1)[1]
    return np.array([int(t) for t in txt.replace("\n", "").split(",") if t.strip()], dtype=np.int32).reshape(512, 2)


def _cvround(v):
    return int(np.rint(v))


class Node:
    __slots__ = ("keys", "x0", "x1", "y0", "y1", "no_more", "seq")

    def __init__(self):
        self.keys = []
        self.no_more = False


def _divide(nd):
    halfx = int(math.ceil(float(f32(nd.x1 - nd.x0) / f32(2))))
    halfy = int(math.ceil(float(f32(nd.y1 - nd.y0) / f32(2))))
    mx, my = nd.x0 + halfx, nd.y0 + halfy
    ch = [Node() for _ in range(4)]
    ch[0].x0, ch[0].x1, ch[0].y0, ch[0].y1 = nd.x0, mx, nd.y0, my
    ch[1].x0, ch[1].x1, ch[1].y0, ch[1].y1 = mx, nd.x1, nd.y0, my
    ch[2].x0, ch[2].x1, ch[2].y0, ch[2].y1 = nd.x0, mx, my, nd.y1
    ch[3].x0, ch[3].x1, ch[3].y0, ch[3].y1 = mx, nd.x1, my, nd.y1
    for k in nd.keys:
        left, top = k[0] < mx, k[1] < my
        ch[(0 if top else 2) if left else (1 if top else 3)].keys.append(k)
    for c in ch:
        if len(c.keys) == 1:
            c.no_more = True
    return ch


def distribute(cands, minX, maxX, minY, maxY, N):
    """cands: list of (x, y, response, idx). Returns idx list in output order."""
    nIni = int(np.round(f32(maxX - minX) / f32(maxY - minY)))
    hX = f32(maxX - minX) / f32(nIni)
    seq = 0
    nodes = []
    for i in range(nIni):
        r = Node()
        r.x0 = int(hX * f32(i)); r.x1 = int(hX * f32(i + 1)); r.y0 = 0; r.y1 = maxY - minY
        r.seq = seq; seq += 1
        nodes.append(r)
    for k in cands:
        nodes[int(f32(k[0]) / hX)].keys.append(k)
    roots = nodes
    nodes = []
    for r in roots:
        if len(r.keys) == 1:
            r.no_more = True
        if r.keys:
            nodes.append(r)
    finished = False
    while not finished:
        prev = len(nodes)
        front, expandable = [], []
        survivors = []
        for nd in nodes:
            if nd.no_more:
                survivors.append(nd)
                continue
            for c in _divide(nd):
                if c.keys:
                    c.seq = seq; seq += 1
                    front.append(c)
                    if len(c.keys) > 1:
                        expandable.append(c)
        nodes = front[::-1] + survivors
        if len(nodes) >= N or len(nodes) == prev:
            finished = True
        elif len(nodes) + 3 * len(expandable) > N:
            while not finished:
                prev = len(nodes)
                order = sorted(expandable, key=lambda n: (len(n.keys), n.seq))
                expandable = []
                front = []
                erased = set()
                size = len(nodes)
                for nd in reversed(order):
                    for c in _divide(nd):
                        if c.keys:
                            c.seq = seq; seq += 1
                            front.append(c)
                            size += 1
                            if len(c.keys) > 1:
                                expandable.append(c)
                    erased.add(id(nd))
                    size -= 1
                    if size >= N:
                        break
                nodes = front[::-1] + [n for n in nodes if id(n) not in erased]
                if len(nodes) >= N or len(nodes) == prev:
                    finished = True
    out = []
    for nd in nodes:
        best = nd.keys[0]
        for k in nd.keys[1:]:
            if k[2] > best[2]:
                best = k
        out.append(best[3])
    return out


class Cv2Extractor:
    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7):
        self.nfeatures, self.nlevels, self.ini, self.min = nfeatures, nlevels, iniThFAST, minThFAST
        s = np.float64(f32(scaleFactor))
        self.sf = [f32(1.0)]
        for i in range(1, nlevels):
            self.sf.append(f32(np.float64(self.sf[-1]) * s))
        self.isf = [f32(1.0) / v for v in self.sf]
        factor = f32(np.float64(1.0) / s)
        want = f32(nfeatures) * (f32(1) - factor) / (f32(1) - f32(math.pow(float(factor), float(nlevels))))
        self.per_level = []
        tot = 0
        for _ in range(nlevels - 1):
            self.per_level.append(_cvround(want)); tot += self.per_level[-1]
            want = f32(want * factor)
        self.per_level.append(max(nfeatures - tot, 0))
        self.umax = [0] * (HALF_PATCH + 1)
        vmax = int(math.floor(float(f32(HALF_PATCH) * np.sqrt(f32(2.0)) / f32(2) + f32(1))))
        vmin = int(math.ceil(float(f32(HALF_PATCH) * np.sqrt(f32(2.0)) / f32(2))))
        for v in range(vmax + 1):
            self.umax[v] = _cvround(math.sqrt(HALF_PATCH * HALF_PATCH - v * v))
        v0 = 0
        for v in range(HALF_PATCH, vmin - 1, -1):
            while self.umax[v0] == self.umax[v0 + 1]:
                v0 += 1
            self.umax[v] = v0
            v0 += 1
        self.pattern = _pattern()
        self.fast_ini = cv2.FastFeatureDetector_create(iniThFAST, True)
        self.fast_min = cv2.FastFeatureDetector_create(minThFAST, True)

    def pyramid(self, img):
        H, W = img.shape
        self.padded, self.levels = [], []
        for l in range(self.nlevels):
            w = _cvround(f32(W) * self.isf[l]); h = _cvround(f32(H) * self.isf[l])
            if l == 0:
                cur = img
            else:
                cur = cv2.resize(self.levels[l - 1], (w, h), interpolation=cv2.INTER_LINEAR)
            pad = cv2.copyMakeBorder(cur, EDGE, EDGE, EDGE, EDGE, cv2.BORDER_REFLECT_101)
            self.padded.append(pad)
            self.levels.append(pad[EDGE:EDGE + h, EDGE:EDGE + w])

    def __call__(self, img):
        self.pyramid(np.ascontiguousarray(img))
        all_kps, all_desc = [], []
        self.candidates = []
        for l in range(self.nlevels):
            lev = self.levels[l]
            h, w = lev.shape
            minB = EDGE - 3
            maxBX, maxBY = w - EDGE + 3, h - EDGE + 3
            width, height = f32(maxBX - minB), f32(maxBY - minB)
            nCols, nRows = int(width / f32(30)), int(height / f32(30))
            wCell = int(math.ceil(float(width / f32(nCols)))); hCell = int(math.ceil(float(height / f32(nRows))))
            cands = []
            for i in range(nRows):
                iniY = minB + i * hCell
                maxY = iniY + hCell + 6
                if iniY >= maxBY - 3:
                    continue
                maxY = min(maxY, maxBY)
                for j in range(nCols):
                    iniX = minB + j * wCell
                    maxX = iniX + wCell + 6
                    if iniX >= maxBX - 6:
                        continue
                    maxX = min(maxX, maxBX)
                    roi = lev[iniY:maxY, iniX:maxX]
                    kc = self.fast_ini.detect(roi)
                    if not kc:
                        kc = self.fast_min.detect(roi)
                    for k in kc:
                        cands.append((int(k.pt[0]) + j * wCell, int(k.pt[1]) + i * hCell, int(k.response), len(cands)))
            self.candidates.append(cands)
            sel = distribute(cands, minB, maxBX, minB, maxBY, self.per_level[l])
            if not sel:
                continue
            blur = cv2.GaussianBlur(lev.copy(), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
            size = float(int(f32(31) * self.sf[l]))
            for idx in sel:
                x, y, r, _ = cands[idx]
                x += minB; y += minB
                ang = self.ic_angle(lev, x, y)
                all_desc.append(self.describe(blur, x, y, ang))
                px, py = f32(x), f32(y)
                if l:
                    px, py = px * self.sf[l], py * self.sf[l]
                all_kps.append((px, py, size, ang, float(r), l, -1))
        return all_kps, (np.stack(all_desc) if all_desc else np.zeros((0, 32), np.uint8))

    def ic_angle(self, lev, x, y):
        m01 = m10 = 0
        for u in range(-HALF_PATCH, HALF_PATCH + 1):
            m10 += u * int(lev[y, x + u])
        for v in range(1, HALF_PATCH + 1):
            d = self.umax[v]
            plus = lev[y + v, x - d:x + d + 1].astype(np.int64)
            minus = lev[y - v, x - d:x + d + 1].astype(np.int64)
            u = np.arange(-d, d + 1)
            m01 += v * int((plus - minus).sum())
            m10 += int((u * (plus + minus)).sum())
        return f32(cv2.fastAtan2(float(m01), float(m10)))

    def describe(self, blur, x, y, ang):
        factorPI = f32(math.pi / 180.0)
        a32 = f32(ang) * factorPI
        # glibc cosf/sinf are correctly rounded in practice; float32(cos(double)) reproduces them except on
        # double-rounding ties, which tests/test_oracle_vs_cv2.py would surface as a descriptor mismatch.
        a = f32(math.cos(float(a32))); b = f32(math.sin(float(a32)))
        px = self.pattern[:, 0].astype(f32); py = self.pattern[:, 1].astype(f32)
        rr = np.rint((px * b).astype(f32) + (py * a).astype(f32)).astype(np.int64)
        cc = np.rint((px * a).astype(f32) - (py * b).astype(f32)).astype(np.int64)
        vals = blur[y + rr, x + cc].astype(np.int32)
        bits = (vals[0::2] < vals[1::2]).astype(np.uint8)
        return np.packbits(bits, bitorder="little")
