"""Randomised parity sweep (tools/fuzz_parity.py): random shapes, extractor parameters and frame kinds, host and device input,
every quadtree kernel form, CUDA against the oracle bit for bit.  A 150 s run of the same sweep (2526 configurations) is how the
kernels of round 1 were last checked; the suite runs it for a few seconds."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_random_configurations_match_the_oracle():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_parity.py"), "8", "3"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "fuzz ok" in r.stdout
