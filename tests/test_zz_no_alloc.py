"""kgcnn_colsum_static: the column sum for callers that cannot bring a device workspace.

Who needs it: an op running inside a framework that has already taken all of HBM for its own arena (a raw cudaMalloc
from the op fails), a C host whose device budget was fixed at start-up, or a serving thread that may not call
cudaMalloc / cudaFree because both synchronise the device.  The entry point keeps its two-stage scratch in a 1 MiB
array that is part of the library image, so the caller passes the input and the output and nothing else.

This file is collected last on purpose (pytest -x stops at the first failure).  It touches nothing but the C-ABI, so
it gives the same answer in a fresh process (where its first call is also the library's first call on the device and
does the address look-up) and after the rest of the suite.
"""
import ctypes

import numpy as np
import pytest
import torch

from gcnn_keras_b200 import _lib

pytestmark = pytest.mark.gpu

MAX_FLOATS = 1 << 18        # KGCNN_COLSUM_STATIC_MAX_FLOATS in include/kgcnn_b200.h
ROWS_PER_BLOCK = 512        # stage-1 rows per block (csrc/activation.cu)
E_STATIC_LIMIT = -5


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _colsum_workspace(a):
    """The existing entry point: the caller allocates the workspace."""
    h = _lib.lib()
    M, U = a.shape
    out = torch.full((U,), float("nan"), dtype=torch.float32, device=a.device)
    nbytes = int(h.kgcnn_colsum_workspace_bytes(M, U))
    ws = torch.empty((max(nbytes, 4),), dtype=torch.uint8, device=a.device)
    _lib.check(h.kgcnn_colsum(_ptr(a), M, U, _ptr(out), _ptr(ws), nbytes, _stream()), "kgcnn_colsum")
    return out


def _colsum_static(a):
    """The new entry point: input and output only."""
    h = _lib.lib()
    M, U = a.shape
    out = torch.full((U,), float("nan"), dtype=torch.float32, device=a.device)
    _lib.check(h.kgcnn_colsum_static(_ptr(a), M, U, _ptr(out), _stream()), "kgcnn_colsum_static")
    return out


def test_static_scratch_colsum_matches_the_workspace_entry_point_twice_in_a_row():
    """Two calls in a row share the array.  The first is large (1 024 stage-1 blocks, half the array), the second
    is smaller with a different width, so it runs over whatever the first one left behind; both must equal the
    workspace version bit for bit (same kernels, same order of additions) and the fp64 sum to fp32 accuracy."""
    torch.cuda.set_device(0)
    rng = np.random.default_rng(2024)
    for M, U in [(524288, 128), (40000, 50)]:
        a_np = rng.standard_normal((M, U)).astype(np.float32)
        a = torch.tensor(a_np).cuda()
        want = _colsum_workspace(a)
        got = _colsum_static(a)
        torch.cuda.synchronize()
        assert torch.equal(got, want), (M, U)
        ref = a_np.astype(np.float64).sum(0)
        err = np.abs(got.cpu().numpy() - ref).max() / np.abs(ref).max()
        assert err < 1e-5, (M, U, err)


def test_static_scratch_colsum_small_and_empty_inputs():
    """One stage-1 block goes straight to the output (no scratch); zero rows give zeros."""
    torch.cuda.set_device(0)
    a = torch.arange(300 * 7, dtype=torch.float32, device="cuda").reshape(300, 7)
    assert torch.equal(_colsum_static(a), _colsum_workspace(a))
    empty = torch.empty((0, 9), dtype=torch.float32, device="cuda")
    assert torch.equal(_colsum_static(empty), torch.zeros(9, device="cuda"))


def test_static_scratch_colsum_rejects_sizes_beyond_the_array():
    """ceil(M / 512) * U above the array size is refused before anything is launched: the output is untouched and
    the error text names the cause.  The largest admissible size still runs."""
    torch.cuda.set_device(0)
    h = _lib.lib()
    U = 128
    M_ok = (MAX_FLOATS // U) * ROWS_PER_BLOCK          # exactly fills the array
    out = torch.full((U,), -7.0, dtype=torch.float32, device="cuda")
    dummy = torch.zeros(1, dtype=torch.float32, device="cuda")
    # one row more needs one more stage-1 block; the size check comes first, so `dummy` is never read
    rc = h.kgcnn_colsum_static(_ptr(dummy), M_ok + 1, U, _ptr(out), _stream())
    assert rc == E_STATIC_LIMIT
    assert b"fixed scratch" in h.kgcnn_error_string(rc)
    torch.cuda.synchronize()
    assert torch.all(out == -7.0)
    with pytest.raises(_lib.KgcnnError, match="fixed scratch"):
        _lib.check(rc, "kgcnn_colsum_static")
    a = torch.ones((M_ok, U), dtype=torch.float32, device="cuda")
    got = _colsum_static(a)
    assert torch.equal(got, _colsum_workspace(a))
    assert torch.all(got == float(M_ok))               # 2^20 ones: exact in fp32
