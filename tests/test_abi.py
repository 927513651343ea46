"""CPU-side checks of the drop-in boundary: liborbx.so loads and exports every symbol include/orbx.h declares
(no compute call is made without a GPU), and the product never touches the oracle."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    txt = open(os.path.join(ROOT, "include", "orbx.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(orb[xm]_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    import orbx
    if not os.path.exists(orbx.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    L = ctypes.CDLL(orbx.LIB_PATH)
    names = _declared()
    assert len(names) >= 35
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert orbx.version().startswith("orbx")
    assert orbx.lib().orbx_last_error() is not None


def test_python_binding_covers_the_abi():
    import orbx
    L = orbx.lib()
    for n in _declared():
        assert getattr(L, n).argtypes is not None or n in ("orbx_last_error", "orbx_version"), n


def test_no_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import orbx
    with pytest.raises(orbx.OrbxError):
        orbx.ORBextractor(1000, 1.2, 8, 20, 7)
    with pytest.raises(orbx.OrbxError):
        orbx.ORBmatcher(0.9, True)


def test_product_does_not_reference_the_oracle():
    pkg = os.path.join(ROOT, "orb_slam2v2-1_b200")
    hits = subprocess.run(["grep", "-rIl", "-e", "oracle_py", "-e", "liborb_oracle", "-e", "orbo_", pkg, os.path.join(ROOT, "include")],
                          capture_output=True, text=True).stdout.split()
    assert hits == [], hits


def test_sincosf_restatement_matches_glibc(tmp_path):
    exe = str(tmp_path / "sc")
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-pthread", "-I", os.path.join(ROOT, "orb_slam2v2-1_b200", "csrc"),
                           "-o", exe, os.path.join(ROOT, "tests", "native", "sincosf_check.cpp")])
    stride = os.environ.get("ORBX_SINCOS_STRIDE", "1")      # 1 = every float in [0, 6.3] (about 4 s on 8 cores)
    n, bs, bc = subprocess.check_output([exe, stride], text=True).split()
    assert int(n) > 1000 and int(bs) == 0 and int(bc) == 0
