"""Handles on a device other than 0 (needs two GPUs; skipped otherwise): extractor shapes whose quadtree needs more than 48 KB of
dynamic shared memory, created on device 0 and device 1 from ONE host thread (the opt-in is per device, not per thread), and a
matcher that follows the calling thread's current device."""
import numpy as np
import pytest

import oracle_py as O
import orbx
from orbx import synth

pytestmark = pytest.mark.gpu


def _two_gpus():
    import torch
    return torch.cuda.device_count() >= 2


@pytest.mark.skipif(not _two_gpus(), reason="needs two CUDA devices")
@pytest.mark.parametrize("w,h,nf,seed", [(1241, 376, 2000, 3000), (640, 480, 1000, 1000)])
def test_extractors_on_two_devices_from_one_thread(w, h, nf, seed):
    img = synth.rects(w, h, seed)
    ok, od = O.OracleExtractor(nf, 1.2, 8, 20, 7)(img)
    for dev in (0, 1, 0):
        ex = orbx.ORBextractor(nf, 1.2, 8, 20, 7, device=dev)
        k, d = ex(img)
        assert k.tobytes() == ok.tobytes() and np.array_equal(d, od), dev


@pytest.mark.skipif(not _two_gpus(), reason="needs two CUDA devices")
def test_matcher_follows_the_current_device():
    import torch
    A, B = synth.descriptor_sets(700, 4700)
    obi, obd, osd = O.match_bruteforce(A, B)
    for dev in (1, 0):
        torch.cuda.set_device(dev)
        m = orbx.ORBmatcher(0.9, True)              # device = -1: the calling thread's current device
        assert m.device() == dev
        bi, bd, sd = m.match_bruteforce(A, B)
        assert np.array_equal(bi, obi) and np.array_equal(bd, obd) and np.array_equal(sd, osd)
        m2 = orbx.ORBmatcher(0.9, True, device=1 - dev)
        assert m2.device() == 1 - dev
        bi, bd, sd = m2.match_bruteforce(A, B)
        assert np.array_equal(bi, obi)
    torch.cuda.set_device(0)
