"""ORBVocabulary::transform (DBoW2 TemplatedVocabulary.h:1140-1270; Frame::ComputeBoW, Frame.cc:410-417).

The reference ships no vocabulary file, so the trees are synthetic (scenes.synthetic_vocabulary).  CPU leg: the C++ oracle against a
pure-Python restatement of the descent and of the BowVector / FeatureVector assembly.  GPU leg: the orbv_* C ABI against the oracle,
bit for bit (word ids, node ids, and the double-precision BowVector values), for every scoring / weighting combination, plus the
text-file loader and the whole chain descriptors -> FeatureVector -> SearchByBoW."""
import numpy as np
import pytest

import oracle_py as O
import scenes
from test_searches import _dist


def py_transform_raw(k, L, parent, desc, weight, feats, levelsup):
    children = [[] for _ in parent]
    for i in range(1, len(parent)):
        children[parent[i]].append(i)
    words = {}
    for i in range(1, len(parent)):
        if not children[i]:
            words[i] = len(words)
    out = []
    for f in feats:
        nid_level = L - levelsup
        nid = 0 if nid_level <= 0 else None
        node, lvl = 0, 0
        while True:
            lvl += 1
            kids = children[node]
            node = min(kids, key=lambda c: (_dist(f, desc[c]), kids.index(c)))
            if lvl == nid_level:
                nid = node
            if not children[node]:
                break
        out.append((words[node], weight[node], node if nid is None else nid))
    return out


def py_transform(raw, scoring, weighting):
    v, fv = {}, {}
    for i, (wid, w, nid) in enumerate(raw):
        if not w > 0:
            continue
        if weighting in (0, 1):
            v[wid] = v.get(wid, 0.0) + w
        else:
            v.setdefault(wid, w)
        fv.setdefault(nid, []).append(i)
    ids = sorted(v)
    vals = [v[i] for i in ids]
    must = scoring != 5
    if weighting in (0, 1) and vals and not must:
        vals = [x / float(len(vals)) for x in vals]
    if must:
        norm = 0.0
        if scoring == 1:
            for x in vals:
                norm += x * x
            norm = float(np.sqrt(norm))
        else:
            for x in vals:
                norm += abs(x)
        if norm > 0:
            vals = [x / norm for x in vals]
    return ids, vals, sorted(fv), [fv[n] for n in sorted(fv)]


@pytest.fixture(scope="module")
def voc():
    (k1, d1), _ = scenes.frames(1000)
    return scenes.synthetic_vocabulary(3, 10, 3, desc_pool=d1), d1


@pytest.mark.parametrize("levelsup", [0, 1, 2, 4])
def test_oracle_transform_raw_vs_python(voc, levelsup):
    (k, L, parent, desc, weight), feats = voc
    ov = O.OracleVocabulary(k, L, parent, desc, weight)
    wid, w, nid = ov.transform_raw(feats[:300], levelsup)
    ref = py_transform_raw(k, L, parent, desc, weight, feats[:300], levelsup)
    assert [(int(a), float(b), int(c)) for a, b, c in zip(wid, w, nid)] == [(a, float(b), c) for a, b, c in ref]
    if levelsup >= L:
        assert (nid == 0).all()
    assert len(set(wid.tolist())) > 50


@pytest.mark.parametrize("scoring,weighting", [(0, 0), (1, 0), (5, 0), (0, 1), (0, 2), (5, 3), (2, 0)])
def test_oracle_transform_vs_python(voc, scoring, weighting):
    (k, L, parent, desc, weight), feats = voc
    ov = O.OracleVocabulary(k, L, parent, desc, weight, scoring, weighting)
    (bid, bval), (fn, fo, fi) = ov.transform(feats, 1)
    ids, vals, nodes, lists = py_transform(py_transform_raw(k, L, parent, desc, weight, feats, 1), scoring, weighting)
    assert bid.tolist() == ids and bval.tolist() == vals
    assert fn.tolist() == nodes and [fi[fo[a]:fo[a + 1]].tolist() for a in range(len(fn))] == lists
    if scoring == 0:
        assert abs(bval.sum() - 1.0) < 1e-12
    assert 0 < len(fi) < len(feats)              # some words are stopped


@pytest.mark.gpu
@pytest.mark.parametrize("k,L,seed", [(10, 3, 3), (10, 4, 4), (7, 5, 5), (40, 2, 6)])
def test_cuda_transform_raw(k, L, seed):
    import orbx
    (k1, d1), (k2, d2) = scenes.frames(1000 + seed)
    kk, LL, parent, desc, weight = scenes.synthetic_vocabulary(seed, k, L, desc_pool=d1)
    v = orbx.ORBVocabulary(kk, LL, parent, desc, weight)
    ov = O.OracleVocabulary(kk, LL, parent, desc, weight)
    feats = np.concatenate([d1, d2, np.random.default_rng(seed).integers(0, 256, (500, 32)).astype(np.uint8)])
    assert v.size() == int((np.bincount(parent[1:], minlength=len(parent)) == 0).sum())
    for levelsup in (0, 1, 2, 4, 9):
        wid, w, nid = v.transform_raw(feats, levelsup)
        owid, ow, onid = ov.transform_raw(feats, levelsup)
        assert np.array_equal(wid, owid) and w.tobytes() == ow.tobytes() and np.array_equal(nid, onid)


@pytest.mark.gpu
@pytest.mark.parametrize("scoring,weighting", [(0, 0), (1, 0), (5, 0), (0, 1), (0, 2), (5, 3), (2, 0), (3, 1), (4, 2)])
def test_cuda_transform_bow_and_feature_vector(scoring, weighting):
    import orbx
    (k1, d1), _ = scenes.frames(1000)
    k, L, parent, desc, weight = scenes.synthetic_vocabulary(3, 10, 4, desc_pool=d1)
    v = orbx.ORBVocabulary(k, L, parent, desc, weight, scoring, weighting)
    ov = O.OracleVocabulary(k, L, parent, desc, weight, scoring, weighting)
    for levelsup in (2, 4):
        (bid, bval), fv = v.transform(d1, levelsup)
        (obid, obval), ofv = ov.transform(d1, levelsup)
        assert np.array_equal(bid, obid) and bval.tobytes() == obval.tobytes()
        for a, b in zip(fv, ofv):
            assert np.array_equal(a, b)
    (bid, bval), fv = v.transform(d1[:0], 4)
    assert len(bid) == 0 and len(fv[0]) == 0 and fv[1].tolist() == [0]


@pytest.mark.gpu
def test_cuda_vocabulary_text_file_and_search_by_bow(tmp_path):
    """descriptors -> transform -> FeatureVector -> SearchByBoW, all through the product, against the same chain on the oracle."""
    import orbx
    (k1, d1), _ = scenes.frames(1000)
    rng = np.random.default_rng(8)
    k2, d2, _ = scenes.second_view(k1, d1, rng, bits=12)
    k, L, parent, desc, weight = scenes.synthetic_vocabulary(9, 10, 4, desc_pool=d1)
    path = str(tmp_path / "voc.txt")
    scenes.write_vocabulary_text(path, k, L, parent, desc, weight)
    v = orbx.ORBVocabulary.loadFromTextFile(path)
    ov = O.OracleVocabulary(k, L, parent, desc, weight)
    _, fv1 = v.transform(d1, 2)
    _, fv2 = v.transform(d2, 2)
    _, ofv1 = ov.transform(d1, 2)
    _, ofv2 = ov.transform(d2, 2)
    for a, b in zip(fv1 + fv2, ofv1 + ofv2):
        assert np.array_equal(a, b)
    valid1 = (rng.random(len(k1)) < 0.9).astype(np.uint8)
    m = orbx.ORBmatcher(0.7, True)
    n, mt = m.SearchByBoW(k1, d1, valid1, fv1, k2, d2, fv2)
    on, omt = O.search_by_bow(k1, d1, valid1, ofv1, k2, d2, ofv2, None, 0.7, True)
    assert n == on and np.array_equal(mt, omt)
    assert n > 100
    with pytest.raises(orbx.OrbxError):
        orbx.ORBVocabulary(10, 3, np.array([0, 0, 5, 1], np.int32), desc[:4], weight[:4])      # parent after its child
