"""CPU: the C-ABI library builds for sm_100a, loads, and exports every symbol include/find_b200.h declares
(no compute calls — there is no GPU here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from find_b200 import build, _lib
    build.build()
    return _lib.load()


def _declared():
    src = open(os.path.join(ROOT, "include", "find_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(findb200_\w+)\s*\(", src)))


def test_header_symbols_exported(lib):
    names = _declared()
    assert len(names) >= 8
    raw = ctypes.CDLL(os.path.join(ROOT, "find_b200", "libfindb200.so"))
    for n in names:
        assert hasattr(raw, n), f"{n} declared in include/find_b200.h but not exported"


def test_python_signatures_cover_header(lib):
    from find_b200 import _lib
    assert sorted(_lib.SIGNATURES) == _declared()


def test_version_and_workspace_query(lib):
    assert b"sm_100a" in lib.findb200_version()
    assert lib.findb200_knn_workspace_bytes(8, 10242, 10000) > 8 * (10242 + 10000) * 8
    assert lib.findb200_knn_workspace_bytes(-1, 1, 1) == 0


def test_sass_is_blackwell_native():
    """cuobjdump proof that the hot kernels use the sm_100a paths (packed fp32x2, bulk copy, tcgen05)."""
    import shutil
    import subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    sass = subprocess.run([cuobjdump, "-sass", os.path.join(ROOT, "find_b200", "libfindb200.so")],
                          capture_output=True, text=True).stdout
    for mnemonic in ("FFMA2", "FADD2", "FMNMX3", "UBLKCP", "UTCHMMA", "LDTM"):
        assert mnemonic in sass, f"{mnemonic} missing from SASS"


def test_no_cpu_fallback():
    import torch
    import find_b200
    with pytest.raises(RuntimeError, match="CUDA"):
        find_b200.chamfer_distance(torch.zeros(1, 4, 3), torch.zeros(1, 4, 3))
