"""GPU: findb200_knn1_static — the 1-NN entry point that allocates nothing (its scratch is a static device array inside the
library, include/find_b200.h) — against findb200_knn1 with a caller-provided workspace, through the C ABI.

The bar is the one of tests/test_chamfer_gpu.py: indices and distances BIT-EXACT.  Two calls in a row (the second on different,
smaller and ragged clouds, so that stale contents of the shared array would show), sizes beyond the limit are rejected with
FINDB200_ERR_WORKSPACE, and the call leaves CUDA's allocator alone (device free memory is unchanged across it).

The file is named so that pytest collects it last: it was added outside the hot-path scope
and must not hide a failure of the parity tests under `pytest -x`.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

ERR_WORKSPACE = -3


def _clouds(seed, n, p1, p2, dev):
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(n, p1, 3, generator=g) * torch.tensor([0.125, 0.05, 0.05])).contiguous().to(dev)
    y = (torch.randn(n, p2, 3, generator=g) * torch.tensor([0.125, 0.05, 0.05])).contiguous().to(dev)
    y[:, : min(p1, p2) // 4] = x[:, : min(p1, p2) // 4]          # exact hits: zero distances and ties on the lowest index
    return x, y


def _both(lib, _lib, x, y, l1, l2):
    """(dists, idx) of findb200_knn1 (torch-allocated workspace) and of findb200_knn1_static on the same inputs."""
    dev = x.device
    n, p1, p2 = x.shape[0], x.shape[1], y.shape[1]
    st = _lib.stream_ptr(dev)
    nb = lib.findb200_knn_workspace_bytes(n, p1, p2)
    ws = torch.empty(nb, dtype=torch.uint8, device=dev)
    d0 = torch.full((n, p1), -1.0, device=dev)
    i0 = torch.full((n, p1), -1, dtype=torch.int64, device=dev)
    d1, i1 = d0.clone(), i0.clone()
    rc = lib.findb200_knn1(_lib.ptr(x), _lib.ptr(y), _lib.ptr(l1), _lib.ptr(l2), n, p1, p2, _lib.ptr(d0), _lib.ptr(i0),
                           _lib.ptr(ws), nb, st)
    _lib.check(rc, "findb200_knn1")
    torch.cuda.synchronize(dev)
    free_before = torch.cuda.mem_get_info(dev)[0]
    rc = lib.findb200_knn1_static(_lib.ptr(x), _lib.ptr(y), _lib.ptr(l1), _lib.ptr(l2), n, p1, p2, _lib.ptr(d1), _lib.ptr(i1), st)
    _lib.check(rc, "findb200_knn1_static")
    torch.cuda.synchronize(dev)
    assert torch.cuda.mem_get_info(dev)[0] == free_before, "findb200_knn1_static changed the device's free memory"
    return (d0, i0), (d1, i1)


def test_knn1_static_matches_knn1_on_two_calls_in_a_row():
    from find_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda:0")
    lib.findb200_set_knn_mode(0, 0)
    # call 1: full clouds near the limit's intended use (4 x 10 242 x 10 000: 3 MiB of scratch)
    x, y = _clouds(5, 4, 10242, 10000, dev)
    (d0, i0), (d1, i1) = _both(lib, _lib, x, y, None, None)
    assert torch.equal(i0, i1)
    assert torch.equal(d0.view(torch.int32), d1.view(torch.int32))
    assert int(i1.min()) >= 0 and int(i1.max()) < 10000 and float(d1.min()) == 0.0
    # call 2, right after: smaller, ragged clouds over the SAME static array (it still holds call 1's sorted clouds and keys)
    x, y = _clouds(6, 3, 777, 1500, dev)
    l1 = torch.tensor([777, 1, 500], dtype=torch.int64, device=dev)
    l2 = torch.tensor([1500, 900, 3], dtype=torch.int64, device=dev)
    (d0, i0), (d1, i1) = _both(lib, _lib, x, y, l1, l2)
    assert torch.equal(i0, i1)
    assert torch.equal(d0.view(torch.int32), d1.view(torch.int32))
    # and the first shapes again: the array is reusable in any order
    x, y = _clouds(7, 4, 10242, 10000, dev)
    (d0, i0), (d1, i1) = _both(lib, _lib, x, y, None, None)
    assert torch.equal(i0, i1) and torch.equal(d0.view(torch.int32), d1.view(torch.int32))


def test_knn1_static_rejects_sizes_beyond_its_limit():
    from find_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda:0")
    n, p1, p2 = 16, 10242, 10000                         # 11.7 MiB of scratch > 8 MiB
    assert lib.findb200_knn_workspace_bytes(n, p1, p2) > (8 << 20)
    x = torch.zeros(1, device=dev)                       # never dereferenced: the size check comes first
    rc = lib.findb200_knn1_static(_lib.ptr(x), _lib.ptr(x), None, None, n, p1, p2, _lib.ptr(x), _lib.ptr(x), _lib.stream_ptr(dev))
    assert rc == ERR_WORKSPACE
    msg = _lib.last_error()
    assert "FINDB200_KNN1_STATIC_MAX_BYTES" in msg and "findb200_knn1" in msg
    torch.cuda.synchronize(dev)
