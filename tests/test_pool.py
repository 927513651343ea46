"""orbx_pool_* (include/orbx.h): the multi-GPU driver inside the library.  Frames are sharded over devices in
chunks, every device runs several host threads with their own handles, results land at the frame's own index: the output bytes must
not depend on the device count, the chunk size or the lane count, and must equal the CPU oracle's."""
import numpy as np
import pytest

import oracle_py as O
import orbx
from orbx import synth


def test_pool_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(orbx.OrbxError):
        orbx.ExtractorPool(1000, devices=[0])


def _frames(n, w=320, h=240, seed=40):
    return [synth.rects(w, h, seed + i) for i in range(n)]


@pytest.mark.gpu
@pytest.mark.parametrize("chunk,lanes", [(1, 1), (3, 2), (4, 3), (16, 2)])
def test_pool_one_device_equals_oracle(chunk, lanes):
    frames = _frames(10) + [synth.flat(320, 240)]           # ragged last chunk, one frame without keypoints
    oe = O.OracleExtractor(500, 1.2, 8, 20, 7)
    want = [oe(f) for f in frames]
    pool = orbx.ExtractorPool(500, 1.2, 8, 20, 7, devices=[0], chunk=chunk, lanes=lanes)
    for _ in range(2):                                       # second call: planned handles, recycled staging
        got = pool.extract_batch(frames)
        assert len(got) == len(want)
        for (k, d), (ok, od) in zip(got, want):
            assert k.tobytes() == ok.tobytes() and np.array_equal(d, od)
    assert len(got[-1][0]) == 0
    pool.close()


@pytest.mark.gpu
def test_pool_output_bytes_do_not_depend_on_the_device_count():
    import torch
    G = torch.cuda.device_count()
    frames = _frames(13, 640, 480, 1000)
    ref = None
    for g in sorted({1, min(2, G), min(4, G), G}):
        pool = orbx.ExtractorPool(1000, devices=list(range(g)), chunk=2, lanes=2)
        assert pool.devices() == g
        got = pool.extract_batch(frames)
        blob = b"".join(k.tobytes() + d.tobytes() for k, d in got)
        if ref is None:
            ref = blob
            ok, od = O.OracleExtractor(1000, 1.2, 8, 20, 7)(frames[5])
            assert got[5][0].tobytes() == ok.tobytes() and np.array_equal(got[5][1], od)
        assert blob == ref, "outputs differ between 1 and %d devices" % g
        pool.close()


@pytest.mark.gpu
def test_pool_reports_errors_of_its_workers():
    import ctypes as C
    L = orbx.lib()
    pool = orbx.ExtractorPool(500, devices=[0], chunk=2, lanes=2)
    frames = _frames(5)
    cap = 8                                                  # far too small: every worker hits ORBX_ERR_CAPACITY
    kps = np.zeros((5, cap), orbx.KP_DTYPE)
    desc = np.zeros((5, cap, 32), np.uint8)
    n = np.zeros(5, np.int32)
    ptrs = (C.c_void_p * 5)(*[f.ctypes.data for f in frames])
    rc = L.orbx_pool_extract_batch(pool._h, ptrs, 5, 320, 240, 320, kps.ctypes.data_as(C.c_void_p), desc.ctypes.data_as(C.c_void_p), cap,
                                   n.ctypes.data_as(C.c_void_p))
    assert rc == orbx.ERR_CAPACITY and b"device 0" in L.orbx_last_error()
    with pytest.raises(orbx.OrbxError):
        orbx.ExtractorPool(500, devices=[0, 0])
    pool.close()


@pytest.mark.gpu
@pytest.mark.parametrize("chunk,lanes", [(2, 3), (5, 2)])
def test_pool_batches_in_flight_equal_synchronous_calls(chunk, lanes):
    """orbx_pool_submit_batch / orbx_pool_wait: several batches queued at once (different sizes, one with a frame without keypoints, one
    empty), waited for out of order -- every batch gets what a synchronous call gives, and an unknown ticket is an error."""
    oe = O.OracleExtractor(500, 1.2, 8, 20, 7)
    batches = [_frames(7, seed=60), _frames(3, seed=80) + [synth.flat(320, 240)], [], _frames(11, seed=100)]
    pool = orbx.ExtractorPool(500, 1.2, 8, 20, 7, devices=[0], chunk=chunk, lanes=lanes)
    for _ in range(2):
        tickets = [pool.submit(b) for b in batches]
        for i in (3, 0, 2, 1):
            got = pool.wait(tickets[i])
            assert len(got) == len(batches[i])
            for (k, d), f in zip(got, batches[i]):
                ok, od = oe(f)
                assert k.tobytes() == ok.tobytes() and np.array_equal(d, od)
    assert orbx.lib().orbx_pool_wait(pool._h, 12345) == orbx.ERR_ARG
    sync = pool.extract_batch(batches[0])                    # the synchronous form still works afterwards
    assert len(sync) == 7
    pool.close()
