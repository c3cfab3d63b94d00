"""Pins the oracle's fixed-radius kNN (oracle/tp_ops.py frnn_grid_points) against the REFERENCE's own CPU
implementation, FRNNBruteForceCPU (third_party/FRNN/frnn/csrc/bruteforce/bruteforce_cpu.cpp), compiled from the
reference tree into oracle/_ref/ by oracle/Makefile.  The reference tests `dist < r2` (strict) and breaks
distance ties by heap order, the grid kernel `<=` (find_nbrs.cu:490): inputs are random floats, for which
neither a tie nor a point exactly at distance r occurs, and the test asserts that precondition."""
import ctypes
import os

import pytest
import torch

from oracle import tp_ops as tp

SO = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libref_frnn_bf.so")
pytestmark = pytest.mark.skipif(not os.path.exists(SO), reason="oracle/_ref not built (needs /root/reference at build time)")


def _ref(p1, p2, l1, l2, K, r):
    lib = ctypes.CDLL(SO)
    N, P1, P2 = p1.shape[0], p1.shape[1], p2.shape[1]
    idx = torch.empty((N, P1, K), dtype=torch.int64)
    d = torch.empty((N, P1, K), dtype=torch.float32)
    vp = ctypes.c_void_p
    lib.ref_frnn_bruteforce_cpu.argtypes = [vp, vp, vp, vp, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_int,
                                            ctypes.c_float, vp, vp]
    rc = lib.ref_frnn_bruteforce_cpu(p1.data_ptr(), p2.data_ptr(), l1.data_ptr(), l2.data_ptr(), N, P1, P2, K, r,
                                     idx.data_ptr(), d.data_ptr())
    assert rc == 0
    return idx, d


@pytest.mark.parametrize("K,r", [(8, 0.1), (20, 0.15), (32, 0.08), (3, 0.5)])
def test_oracle_frnn_equals_reference_bruteforce(K, r):
    g = torch.Generator().manual_seed(100 + K)
    N, P1, P2 = 2, 300, 500
    p1 = torch.rand(N, P1, 3, generator=g).contiguous()
    p2 = torch.rand(N, P2, 3, generator=g).contiguous()
    l1 = torch.tensor([P1, P1 - 37], dtype=torch.int64)
    l2 = torch.tensor([P2 - 11, P2], dtype=torch.int64)
    ridx, rd = _ref(p1, p2, l1, l2, K, r)
    out = tp.frnn_grid_points(p1, p2, l1, l2, K, r)
    d, idx = out[0], out[1]
    for n in range(N):
        a, b = ridx[n, :l1[n]], idx[n, :l1[n]]
        da = rd[n, :l1[n]]
        # precondition: no distance ties inside a row, nothing exactly on the sphere
        valid = a >= 0
        srt = torch.where(valid, da, torch.full_like(da, float("inf")))
        assert not ((srt[:, 1:] == srt[:, :-1]) & valid[:, 1:]).any()
        assert torch.equal(a, b)
        assert torch.equal(torch.where(valid, da, torch.zeros_like(da)),
                           torch.where(valid, d[n, :l1[n]], torch.zeros_like(da)))


def test_self_query_contains_self_first():
    g = torch.Generator().manual_seed(5)
    p = torch.rand(1, 400, 3, generator=g)
    l = torch.tensor([400])
    ridx, rd = _ref(p, p, l, l, 4, 0.2)
    assert torch.equal(ridx[0, :, 0], torch.arange(400))
    assert torch.equal(tp.frnn_grid_points(p, p, l, l, 4, 0.2)[1][0], ridx[0])
