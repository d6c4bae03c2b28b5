"""GPU: tcgen05 self-test — validates the swizzled tile-image layout, the K-major / MN-major shared-memory
descriptors, the instruction descriptor and the TMEM lane/column mapping that csrc/field.cu builds on."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def to_image(mat: np.ndarray) -> np.ndarray:
    """[rows, 64*nblk] fp16 -> nblk images of [rows x 64] (row r at r*128 B, 16-B chunk c at c ^ (r & 7))."""
    rows, cols = mat.shape
    assert cols % 64 == 0
    out = np.zeros((cols // 64, rows, 8, 8), np.float16)
    for cb in range(cols // 64):
        blk = mat[:, cb * 64:(cb + 1) * 64].reshape(rows, 8, 8)
        for r in range(rows):
            out[cb, r, np.arange(8) ^ (r & 7)] = blk[r]
    return out.reshape(-1)


def _run(mode, a_img, b_img):
    from find_b200 import _lib
    lib = _lib.load()
    a = torch.from_numpy(a_img.view(np.int16)).cuda()
    b = torch.from_numpy(b_img.view(np.int16)).cuda()
    d = torch.zeros(128, 256, device="cuda")
    rc = lib.findb200_umma_selftest(mode, _lib.ptr(a), _lib.ptr(b), _lib.ptr(d), _lib.stream_ptr(d.device))
    _lib.check(rc, "findb200_umma_selftest")
    torch.cuda.synchronize()
    return d.cpu().numpy()


def test_kmajor_forward_shape():
    rng = np.random.RandomState(0)
    A = (rng.randn(128, 256) * 0.5).astype(np.float16)
    W = (rng.randn(256, 256) * 0.1).astype(np.float16)      # [N, K] like nn.Linear.weight
    b_img = np.concatenate([to_image(W[nh * 128:(nh + 1) * 128, kc * 64:(kc + 1) * 64])
                            for nh in range(2) for kc in range(4)])
    D = _run(0, to_image(A), b_img)
    ref = A.astype(np.float32) @ W.astype(np.float32).T
    np.testing.assert_allclose(D, ref, rtol=1e-3, atol=1e-3)


def test_mnmajor_dw_shape():
    rng = np.random.RandomState(1)
    G = (rng.randn(128, 128) * 0.5).astype(np.float16)      # [rows, m]
    H = (rng.randn(128, 256) * 0.5).astype(np.float16)      # [rows, n]
    D = _run(1, to_image(G), to_image(H))
    ref = G.astype(np.float32).T @ H.astype(np.float32)
    np.testing.assert_allclose(D, ref, rtol=1e-3, atol=2e-3)


def _run_pair(mode, A, W):
    """CTA-pair self-test (modes 2 / 3): A [256,256], W [256,256] fp16 -> D [256,256] through tcgen05.mma.cta_group::2."""
    from find_b200 import _lib
    lib = _lib.load()
    a_img = np.concatenate([to_image(A[r * 128:(r + 1) * 128]) for r in range(2)])
    b_img = np.concatenate([to_image(W[:, kc * 64:(kc + 1) * 64]) for kc in range(4)])
    a = torch.from_numpy(a_img.view(np.int16)).cuda()
    b = torch.from_numpy(b_img.view(np.int16)).cuda()
    d = torch.zeros(256, 256, device="cuda")
    rc = lib.findb200_umma_selftest(mode, _lib.ptr(a), _lib.ptr(b), _lib.ptr(d), _lib.stream_ptr(d.device))
    _lib.check(rc, "findb200_umma_selftest")
    torch.cuda.synchronize()
    return d.cpu().numpy()


def test_cta_pair_m256_relay():
    """cta_group::2: each CTA stages its 128 rows of A and HALF of every weight chunk; peer -> leader relay."""
    rng = np.random.RandomState(2)
    A = (rng.randn(256, 256) * 0.5).astype(np.float16)
    W = (rng.randn(256, 256) * 0.1).astype(np.float16)
    D = _run_pair(2, A, W)
    ref = A.astype(np.float32) @ W.astype(np.float32).T
    np.testing.assert_allclose(D, ref, rtol=1e-3, atol=1e-3)
