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
