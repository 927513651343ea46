"""Golden fixtures (tests/golden/*.npz, written by tools/gen_golden.py).

The reference ships no golden vectors, so the fixtures were produced by the C++ oracle after
it agreed byte for byte with an independent Python orchestration over real cv2 4.13 primitives.  CPU leg: the oracle
built in this checkout still reproduces them (guards against drift of the checker).  GPU leg: the CUDA path through
the C ABI reproduces them without the oracle in the loop."""
import glob
import os

import numpy as np
import pytest

import oracle_py as O
from orbx import synth

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
EXTRACT = sorted(glob.glob(os.path.join(GOLD, "extract_*.npz")))
KPF = ["x", "y", "size", "angle", "response", "octave", "class_id"]


def _ids(paths):
    return [os.path.basename(p)[:-4] for p in paths]


def test_fixtures_present():
    assert len(EXTRACT) >= 5
    assert os.path.exists(os.path.join(GOLD, "match_bruteforce_2000_s6000.npz"))


def _check_extract(g, kps, desc, level_counts, cand_counts, cand_l0, cand_l7, level7, blurred7):
    assert len(kps) == len(g["kps"])
    for f in KPF:
        assert np.array_equal(kps[f], g["kps"][f]), f
    assert np.array_equal(desc, g["desc"])
    assert list(level_counts) == list(g["level_counts"])
    assert list(cand_counts) == list(g["cand_counts"])
    assert np.array_equal(cand_l0, g["cand_l0"]) and np.array_equal(cand_l7, g["cand_l7"])
    assert np.array_equal(level7, g["level7"])
    assert np.array_equal(blurred7, g["blurred7"])


@pytest.mark.parametrize("path", EXTRACT, ids=_ids(EXTRACT))
def test_oracle_reproduces_golden(path):
    g = np.load(path)
    img = synth.frame(str(g["kind"]), int(g["w"]), int(g["h"]), int(g["seed"]))
    oe = O.OracleExtractor(int(g["nf"]), 1.2, 8, 20, 7)
    kps, desc = oe(img)
    cand = [oe.level_candidates(l) for l in range(8)]
    _check_extract(g, kps, desc, [oe.level_nkeys(l) for l in range(8)], [len(c) for c in cand], cand[0], cand[7],
                   oe.level_padded(7), oe.level_blurred(7))


def test_oracle_matcher_reproduces_golden():
    g = np.load(os.path.join(GOLD, "match_bruteforce_2000_s6000.npz"))
    A, B = synth.descriptor_sets(int(g["N"]), int(g["seed"]))
    bi, bd, sd = O.match_bruteforce(A, B)
    assert np.array_equal(bi, g["best_idx"]) and np.array_equal(bd, g["best_dist"]) and np.array_equal(sd, g["second_dist"])


@pytest.mark.gpu
@pytest.mark.parametrize("path", EXTRACT, ids=_ids(EXTRACT))
def test_cuda_reproduces_golden(path):
    import orbx
    g = np.load(path)
    img = synth.frame(str(g["kind"]), int(g["w"]), int(g["h"]), int(g["seed"]))
    ex = orbx.ORBextractor(int(g["nf"]), 1.2, 8, 20, 7)
    kps, desc = ex(img)
    cand = [ex.level_candidates(l) for l in range(8)]
    _check_extract(g, kps, desc, [ex.level_count(l) for l in range(8)], [len(c) for c in cand], cand[0], cand[7],
                   ex.pyramid_level(7, 0, border=19), ex.blurred_level(7))


@pytest.mark.gpu
def test_cuda_matcher_reproduces_golden():
    import orbx
    g = np.load(os.path.join(GOLD, "match_bruteforce_2000_s6000.npz"))
    A, B = synth.descriptor_sets(int(g["N"]), int(g["seed"]))
    bi, bd, sd = orbx.ORBmatcher(0.9, True).match_bruteforce(A, B)
    assert np.array_equal(bi, g["best_idx"]) and np.array_equal(bd, g["best_dist"]) and np.array_equal(sd, g["second_dist"])
