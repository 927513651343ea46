"""orbm_match_bruteforce_device_noalloc: the brute-force scan with its scratch in a fixed file-scope device array (no cudaMalloc on
any path; include/orbx.h).  It has no counterpart in the reference.

Checked through the C ABI against the existing handle-based entry point orbm_match_bruteforce_device on the same device buffers,
bit for bit (integer work: the tolerance of the matcher tests is zero), on two calls in a row with different problem sizes -- the
array is shared and keeps the last call's contents.  This file sorts last on purpose."""
import ctypes as C

import numpy as np
import pytest

import orbx
from orbx import synth

pytestmark = pytest.mark.gpu


def _both(L, m, A, B, stream):
    import torch
    nq, nt = len(A), len(B)
    dq, dt = torch.from_numpy(A).cuda(), torch.from_numpy(B).cuda()
    ref = [torch.full((max(nq, 1),), -7, dtype=torch.int32, device="cuda") for _ in range(3)]
    got = [torch.full((max(nq, 1),), -9, dtype=torch.int32, device="cuda") for _ in range(3)]
    torch.cuda.synchronize()
    rc = L.orbm_match_bruteforce_device(m._h, dq.data_ptr(), nq, dt.data_ptr(), nt, *[x.data_ptr() for x in ref])
    assert rc == 0, L.orbx_last_error()
    rc = L.orbm_match_bruteforce_device_noalloc(dq.data_ptr(), nq, dt.data_ptr(), nt, *[x.data_ptr() for x in got], C.c_void_p(stream.cuda_stream))
    assert rc == 0, L.orbx_last_error()
    stream.synchronize()
    torch.cuda.synchronize()
    return [x.cpu().numpy()[:nq] for x in ref], [x.cpu().numpy()[:nq] for x in got]


def test_noalloc_equals_handle_based_entry_two_calls_in_a_row():
    import torch
    L = orbx.lib()
    m = orbx.ORBmatcher(0.9, True)
    stream = torch.cuda.Stream()
    # call 1: many segments per query block; call 2, straight after on the same stream: other sizes, a ragged last block, and a
    # train set small enough for a single segment per block
    for nq, nt, seed in [(3000, 2500, 4100), (1111, 777, 4200), (130, 40, 4300)]:
        A, _ = synth.descriptor_sets(nq, seed)
        _, B = synth.descriptor_sets(nt, seed + 1)
        B[: min(nq, nt) // 2] = A[: min(nq, nt) // 2]          # exact matches and ties among them
        ref, got = _both(L, m, np.ascontiguousarray(A), np.ascontiguousarray(B), stream)
        for r, g, name in zip(ref, got, ("best_idx", "best_dist", "second_dist")):
            assert np.array_equal(r, g), (nq, nt, name)
        assert (ref[1][: min(nq, nt) // 2] == 0).all()


def test_noalloc_maximum_size_and_rejection():
    import torch
    L = orbx.lib()
    m = orbx.ORBmatcher(0.9, True)
    stream = torch.cuda.Stream()
    A, B = synth.descriptor_sets(1024, 4400)
    A = np.ascontiguousarray(np.tile(A, (16, 1)))             # 16384 queries: the stated maximum
    ref, got = _both(L, m, A, np.ascontiguousarray(B), stream)
    for r, g in zip(ref, got):
        assert np.array_equal(r, g)
    big = torch.zeros((16385, 32), dtype=torch.uint8, device="cuda")
    o = [torch.zeros(16385, dtype=torch.int32, device="cuda") for _ in range(3)]
    rc = L.orbm_match_bruteforce_device_noalloc(big.data_ptr(), 16385, big.data_ptr(), 16, *[x.data_ptr() for x in o], C.c_void_p(0))
    assert rc == orbx.ERR_ARG and b"beyond the fixed scratch" in L.orbx_last_error()
    # empty problems are accepted and touch nothing
    assert L.orbm_match_bruteforce_device_noalloc(None, 0, None, 0, None, None, None, C.c_void_p(0)) == 0
