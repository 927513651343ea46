"""Parity of the CUDA Hamming matcher with the CPU restatement of ORBmatcher's scans: indices bit-exact."""
import numpy as np
import pytest

import oracle_py as O
import orbx
from orbx import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def m():
    return orbx.ORBmatcher(0.9, True)


def test_descriptor_distance(m):
    A, B = synth.descriptor_sets(500, 4001)
    d = m.DescriptorDistance(A, B)
    ref = np.array([O.descriptor_distance(a, b) for a, b in zip(A, B)])
    assert np.array_equal(d, ref)
    assert m.DescriptorDistance(A[0], A[0]) == 0
    assert m.DescriptorDistance(np.zeros(32, np.uint8), np.full(32, 255, np.uint8)) == 256


@pytest.mark.parametrize("N", [2000, 10000])
def test_bruteforce_config4(m, N):
    A, B = synth.descriptor_sets(N, 4000 + N)
    bi, bd, sd = m.match_bruteforce(A, B)
    obi, obd, osd = O.match_bruteforce(A, B)
    assert np.array_equal(bi, obi) and np.array_equal(bd, obd) and np.array_equal(sd, osd)
    for ratio in (0.9, 0.7):
        _, _, _, acc = m.match_ratio(A, B, 50, ratio)
        ref = (obd <= 50) & (obd.astype(np.float32) < np.float32(ratio) * osd.astype(np.float32))
        assert np.array_equal(acc, ref)
        assert acc.sum() > N // 3


def test_bruteforce_ragged_and_ties(m):
    rng = np.random.default_rng(3)
    for nq, nt in [(1, 1), (7, 3), (130, 1), (1, 700), (129, 257), (300, 65), (5, 0)]:
        q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
        t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
        if nt > 4:
            t[nt // 2] = t[0]; t[nt - 1] = t[1]          # equal distances: lowest index must win
            q[0] = t[0]
        bi, bd, sd = m.match_bruteforce(q, t)
        if nt == 0:
            assert (bi == -1).all() and (bd == 256).all() and (sd == 256).all()
            continue
        obi, obd, osd = O.match_bruteforce(q, t)
        assert np.array_equal(bi, obi) and np.array_equal(bd, obd) and np.array_equal(sd, osd)
    # few distinct rows -> many ties; complements -> distance 256 must never be selected
    t = np.repeat(rng.integers(0, 256, (4, 32), dtype=np.uint8), 50, axis=0)
    q = np.concatenate([t[::7], 255 - t[:3]])
    bi, bd, sd = m.match_bruteforce(q, np.concatenate([t, 255 - q[:1]]))
    obi, obd, osd = O.match_bruteforce(q, np.concatenate([t, 255 - q[:1]]))
    assert np.array_equal(bi, obi) and np.array_equal(bd, obd) and np.array_equal(sd, osd)
    one = rng.integers(0, 256, (1, 32), dtype=np.uint8)
    bi, bd, sd = m.match_bruteforce(one, 255 - one)
    assert bi[0] == -1 and bd[0] == 256 and sd[0] == 256
    assert O.match_bruteforce(one, 255 - one)[0][0] == -1


def test_candidate_lists(m):
    rng = np.random.default_rng(5)
    A, B = synth.descriptor_sets(2000, 6000)
    lens = rng.integers(0, 120, len(A))
    lens[:5] = [0, 1, 2, 33, 64]
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    idx = rng.integers(0, len(B), off[-1]).astype(np.int32)     # repeats allowed
    bi, bd, si, sd = m.match_candidates(A, B, off, idx)
    obi, obd, osi, osd = O.match_candidates(A, B, off, idx)
    assert np.array_equal(bi, obi) and np.array_equal(bd, obd) and np.array_equal(si, osi) and np.array_equal(sd, osd)
    d = m.candidate_distances(A, B, off, idx)
    q_of = np.repeat(np.arange(len(A)), lens)
    ref = np.unpackbits(A[q_of] ^ B[idx], axis=1).sum(axis=1)
    assert np.array_equal(d, ref)


def test_bruteforce_device_resident(m):
    import torch
    A, B = synth.descriptor_sets(2000, 6000)
    dq, dt = torch.from_numpy(A).cuda(), torch.from_numpy(B).cuda()
    out = [torch.zeros(len(A), dtype=torch.int32, device="cuda") for _ in range(3)]
    torch.cuda.synchronize()
    m.match_bruteforce_device(dq.data_ptr(), len(A), dt.data_ptr(), len(B), *[o.data_ptr() for o in out])
    torch.cuda.current_stream().synchronize()
    s = torch.cuda.ExternalStream(m.stream())
    s.synchronize()
    obi, obd, osd = O.match_bruteforce(A, B)
    assert np.array_equal(out[0].cpu().numpy(), obi) and np.array_equal(out[1].cpu().numpy(), obd) and np.array_equal(out[2].cpu().numpy(), osd)
