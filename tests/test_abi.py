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


def test_product_path_never_touches_the_oracle_and_fails_loudly_without_cuda():
    """The oracle is test infrastructure: nothing under find_b200/ may import or execute it; CPU tensors raise instead of
    falling back (the renderer, the mesh ops and the step all go through _lib.require_cuda)."""
    import os
    import re
    import pytest
    import torch
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pkg = os.path.join(root, "find_b200")
    for dirpath, _, files in os.walk(pkg):
        for name in files:
            if name.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, name), encoding="utf-8").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, re.M), f"{name} imports the oracle"
                assert "oracle/_ref" not in text and "/root/reference/" not in text.replace("/root/reference/src", "").replace("/root/reference/cfgs", ""), name
    from find_b200.renderer import rasterize_silhouette, rasterize_nearest
    from find_b200.mesh import Meshes, sample_points_from_meshes
    v = torch.zeros(1, 3, 3); f = torch.tensor([[0, 1, 2]])
    for call in (lambda: rasterize_silhouette(v, f, 8), lambda: rasterize_nearest(v, f, 8),
                 lambda: sample_points_from_meshes(Meshes(v, f), 4)):
        with pytest.raises(RuntimeError):
            call()
