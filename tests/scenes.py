"""Synthetic tracking scenes for the matcher tests: frames with real ORB keypoints (CPU oracle extractor on the seeded
'rects' generator) plus the per-MapPoint / per-KeyFrame arrays the reference's searches read from its SLAM objects,
drawn from a seeded PRNG so that the oracle and the CUDA path see identical inputs."""
import numpy as np

import oracle_py as O
from orbx import synth

W, H = 640, 480
BOUNDS = (0.0, float(W), 0.0, float(H))
SF = np.array([1.2 ** l for l in range(8)], np.float32)


def frames(seed, nf=1000):
    """(kL, dL), (kR, dR): keypoints/descriptors of a synthetic stereo pair (same scene, shifted rectangles)."""
    L, R = synth.stereo_rects(W, H, seed)[:2]
    oe = O.OracleExtractor(nf, 1.2, 8, 20, 7)
    a = oe(L)
    b = oe(R)
    return a, b


def flip_bits(desc, rng, kmax):
    out = desc.copy()
    for i in range(len(out)):
        for b in rng.integers(0, 256, size=int(rng.integers(0, kmax + 1))):
            out[i, b >> 3] ^= np.uint8(1 << (b & 7))
    return out


def projection_scene(seed, n_random=150):
    """Current frame = left image; 'map points' = its own keypoints seen again with noisy position, a few flipped
    descriptor bits and a neighbouring scale level, plus random distractors."""
    rng = np.random.default_rng(seed)
    (k, d), _ = frames(seed)
    n = len(k)
    sel = rng.permutation(n)[: (2 * n) // 3]
    xy = np.stack([k["x"][sel], k["y"][sel]], 1) + rng.uniform(-3, 3, size=(len(sel), 2))
    lvl = np.clip(k["octave"][sel] + rng.integers(-1, 2, size=len(sel)), 0, 7)
    md = flip_bits(d[sel], rng, 30)
    ang = (k["angle"][sel] + rng.normal(0, 3, size=len(sel))) % 360
    rxy = np.stack([rng.uniform(0, W, n_random), rng.uniform(0, H, n_random)], 1)
    xy = np.concatenate([xy, rxy]).astype(np.float32)
    lvl = np.concatenate([lvl, rng.integers(0, 8, n_random)]).astype(np.int32)
    md = np.concatenate([md, rng.integers(0, 256, size=(n_random, 32)).astype(np.uint8)])
    ang = np.concatenate([ang, rng.uniform(0, 360, n_random)]).astype(np.float32)
    order = rng.permutation(len(xy))
    xy, lvl, md, ang = xy[order], lvl[order], md[order], ang[order]
    nmp = len(xy)
    u_right = np.where(rng.random(n) < 0.4, -1.0, k["x"] - rng.uniform(1, 40, n)).astype(np.float32)
    return dict(kps=k, desc=d, u_right=u_right, occupied=(rng.random(n) < 0.1).astype(np.uint8),
                proj_xy=xy, proj_xr=(xy[:, 0] - rng.uniform(1, 40, nmp)).astype(np.float32), level=lvl,
                view_cos=rng.uniform(0.99, 1.0, nmp).astype(np.float32), mp_desc=md,
                mp_has_obs=(rng.random(nmp) < 0.7).astype(np.uint8), angle=ang)


def feature_vector(desc, valid_nodes=32):
    """CSR stand-in for a DBoW2::FeatureVector: node id = top bits of the first descriptor byte."""
    node_of = (desc[:, 0].astype(np.int32) * valid_nodes) >> 8
    nodes = np.unique(node_of)
    off, idx = [0], []
    for nd in nodes:
        members = np.flatnonzero(node_of == nd)
        idx.append(members)
        off.append(off[-1] + len(members))
    return nodes.astype(np.int32), np.asarray(off, np.int32), np.concatenate(idx).astype(np.int32)


def bow_scene(seed):
    rng = np.random.default_rng(seed + 77)
    (k1, d1), (k2, d2) = frames(seed)
    # second view of the first frame's features: same rows, a few flipped bits, shuffled -> real matches share most nodes
    perm = rng.permutation(len(k1))
    k3, d3 = k1[perm].copy(), flip_bits(d1[perm], rng, 20)
    k3["angle"] = (k3["angle"] + rng.normal(0, 4, len(k3)).astype(np.float32)) % 360
    return dict(k1=k1, d1=d1, valid1=(rng.random(len(k1)) < 0.8).astype(np.uint8), fv1=feature_vector(d1),
                k2=k3, d2=d3, valid2=(rng.random(len(k3)) < 0.8).astype(np.uint8), fv2=feature_vector(d3),
                kR=k2, dR=d2, validR=(rng.random(len(k2)) < 0.8).astype(np.uint8), fvR=feature_vector(d2))


SIGMA2 = (SF * SF).astype(np.float32)
INV_SIGMA2 = (np.float32(1.0) / SIGMA2).astype(np.float32)


def fuse_scene(seed, n_random=200):
    """Key frame = left image; MapPoints of neighbouring key frames projected into it: most land within a pixel or two of
    one of its features (so the chi-square gate of Fuse decides), the rest anywhere."""
    rng = np.random.default_rng(seed + 31)
    (k, d), _ = frames(seed)
    n = len(k)
    sel = rng.permutation(n)[: (3 * n) // 4]
    xy = np.stack([k["x"][sel], k["y"][sel]], 1) + rng.normal(0, 1.2, size=(len(sel), 2)) * SF[k["octave"][sel]][:, None]
    lvl = np.clip(k["octave"][sel] + rng.integers(0, 2, size=len(sel)), 0, 7)
    md = flip_bits(d[sel], rng, 40)
    u_right = np.where(rng.random(n) < 0.4, -1.0, k["x"] - rng.uniform(1, 40, n)).astype(np.float32)
    ur = np.where(u_right[sel] >= 0, u_right[sel] + rng.normal(0, 1.5, len(sel)), xy[:, 0] - rng.uniform(1, 40, len(sel)))
    rxy = np.stack([rng.uniform(0, W, n_random), rng.uniform(0, H, n_random)], 1)
    xy = np.concatenate([xy, rxy]).astype(np.float32)
    lvl = np.concatenate([lvl, rng.integers(0, 8, n_random)]).astype(np.int32)
    md = np.concatenate([md, rng.integers(0, 256, size=(n_random, 32)).astype(np.uint8)])
    ur = np.concatenate([ur, rxy[:, 0] - rng.uniform(1, 40, n_random)]).astype(np.float32)
    order = rng.permutation(len(xy))
    return dict(kps=k, desc=d, u_right=u_right, proj_xy=xy[order], proj_ur=ur[order], level=lvl[order], mp_desc=md[order],
                skip=(rng.random(n) < 0.15).astype(np.uint8))


def second_view(k1, d1, rng, jitter=1.0, max_shift=40.0, bits=20):
    """Same features seen from a camera translated along x: shuffled rows, shifted left by a per-feature disparity, small
    vertical jitter, a few flipped descriptor bits.  Returns (k2, d2, perm) with row j of the second view = row perm[j] of the first."""
    perm = rng.permutation(len(k1))
    k2, d2 = k1[perm].copy(), flip_bits(d1[perm], rng, bits)
    k2["x"] = (k2["x"] - rng.uniform(0, max_shift, len(k2))).astype(np.float32)
    k2["y"] = (k2["y"] + rng.normal(0, jitter, len(k2)) * SF[k2["octave"]]).astype(np.float32)
    k2["angle"] = ((k2["angle"] + rng.normal(0, 4, len(k2))) % 360).astype(np.float32)
    keep = (k2["x"] >= 0) & (k2["y"] >= 0) & (k2["y"] < H)
    return k2[keep].copy(), d2[keep].copy(), perm[keep]


def sim3_scene(seed):
    rng = np.random.default_rng(seed + 53)
    (k1, d1), _ = frames(seed)
    k2, d2, perm = second_view(k1, d1, rng)
    inv = np.full(len(k1), -1)
    inv[perm] = np.arange(len(k2))
    # MapPoints of key frame 1 projected into key frame 2: near their true counterpart when it exists
    has = inv >= 0
    p12 = np.stack([rng.uniform(0, W, len(k1)), rng.uniform(0, H, len(k1))], 1)
    p12[has] = np.stack([k2["x"][inv[has]], k2["y"][inv[has]]], 1) + rng.uniform(-3, 3, size=(int(has.sum()), 2))
    l12 = np.where(has, np.clip(k2["octave"][np.maximum(inv, 0)] + rng.integers(0, 2, len(k1)), 0, 7), rng.integers(0, 8, len(k1)))
    l12 = np.where(rng.random(len(k1)) < 0.2, -1, l12).astype(np.int32)
    p21 = np.stack([k1["x"][perm], k1["y"][perm]], 1) + rng.uniform(-3, 3, size=(len(k2), 2))
    l21 = np.clip(k1["octave"][perm] + rng.integers(0, 2, len(k2)), 0, 7)
    l21 = np.where(rng.random(len(k2)) < 0.2, -1, l21).astype(np.int32)
    return dict(k1=k1, d1=d1, k2=k2, d2=d2, proj12_xy=p12.astype(np.float32), level12=l12, mp1_desc=flip_bits(d1, rng, 15),
                proj21_xy=p21.astype(np.float32), level21=l21, mp2_desc=flip_bits(d2, rng, 15))


def triangulation_scene(seed):
    rng = np.random.default_rng(seed + 91)
    (k1, d1), _ = frames(seed)
    k2, d2, _ = second_view(k1, d1, rng, jitter=1.2)
    # fundamental matrix of a pure x translation (epipolar lines = image rows) with a small perturbation
    F = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float64) + rng.normal(0, 2e-6, (3, 3)) * np.array([[1, 1, 300], [1, 1, 300], [300, 300, 9e4]])
    ur = lambda n: np.where(rng.random(n) < 0.5, -1.0, rng.uniform(1, 600, n)).astype(np.float32)
    return dict(k1=k1, d1=d1, has_mp1=(rng.random(len(k1)) < 0.3).astype(np.uint8), u_right1=ur(len(k1)), fv1=feature_vector(d1, 16),
                k2=k2, d2=d2, has_mp2=(rng.random(len(k2)) < 0.3).astype(np.uint8), u_right2=ur(len(k2)), fv2=feature_vector(d2, 16),
                F12=F.astype(np.float32), epipole=(np.float32(300.25), np.float32(200.5)))


def synthetic_vocabulary(seed, k=10, L=3, shallow_leaf_prob=0.08, desc_pool=None):
    """A k-ary vocabulary tree numbered the way DBoW2 saves it (saveToTextFile walks nodes in id order; ids were assigned in
    breadth-first creation order): children cluster around their parent (a few flipped bits), a few branches stop early (leaves
    above level L, as k-means leaves them in ORBvoc.txt), two siblings per node are identical (distance ties), a few words are
    stopped (weight 0).  Returns (k, L, parent, desc, weight)."""
    rng = np.random.default_rng(seed + 17)
    parent, desc, weight, level = [0], [np.zeros(32, np.uint8)], [0.0], [0]
    frontier = [0]
    for lv in range(1, L + 1):
        nxt = []
        for p in frontier:
            base = desc_pool[rng.integers(0, len(desc_pool))] if (lv == 1 and desc_pool is not None) else desc[p]
            kids = flip_bits(np.repeat(base[None], k, 0), rng, 90 if lv == 1 else 40)
            kids[rng.integers(0, k)] = kids[rng.integers(0, k)]
            for c in range(k):
                parent.append(p); desc.append(kids[c]); weight.append(0.0); level.append(lv)
                if lv < L and rng.random() > shallow_leaf_prob:
                    nxt.append(len(parent) - 1)
        frontier = nxt
    parent = np.asarray(parent, np.int32)
    has_child = np.zeros(len(parent), bool)
    has_child[parent[1:]] = True
    weight = np.where(has_child, 0.0, rng.uniform(0.5, 9.0, len(parent)))
    weight[(~has_child) & (rng.random(len(parent)) < 0.03)] = 0.0
    weight[0] = 0.0
    return k, L, parent, np.stack(desc).astype(np.uint8), weight.astype(np.float64)


def write_vocabulary_text(path, k, L, parent, desc, weight, scoring=0, weighting=0):
    """saveToTextFile (TemplatedVocabulary.h:1441-1468)."""
    has_child = np.zeros(len(parent), bool)
    has_child[parent[1:]] = True
    with open(path, "w") as f:
        f.write("%d %d  %d %d\n" % (k, L, scoring, weighting))
        for i in range(1, len(parent)):
            f.write("%d %d %s %r\n" % (parent[i], 0 if has_child[i] else 1, " ".join(str(int(b)) for b in desc[i]), float(weight[i])))
