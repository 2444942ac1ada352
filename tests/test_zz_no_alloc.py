"""dbaf_ba_tables_noalloc: the graph tables for callers that cannot allocate device memory around
the call (a real-time tracking thread with a fixed memory budget, a plugin inside a host
application that owns all buffers).  Its intermediate tables sit in a fixed array inside the
library, shared by every call, so the test runs it twice in a row on different graphs -- the
second, smaller graph would expose anything left over from the first -- and compares each result
with the tables the regular entry point (dbaf_ba_prepare, via droid_backends.ba_tables) builds in a
caller-allocated workspace.  Graph bookkeeping is integer work: the comparison is exact.

This file is named so that pytest collects it last.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _noalloc_tables(lib, ii, jj, t0, t1, n_frames, dev):
    E, P = ii.numel(), t1 - t0
    kcap = min(n_frames, P + E)
    out = dict(info=torch.full((16,), -7, dtype=torch.int32, device=dev),
               kx=torch.full((kcap,), -7, dtype=torch.int32, device=dev),
               kk_exp=torch.full((P + E,), -7, dtype=torch.int32, device=dev),
               csr_ptr=torch.full((kcap + 1,), -7, dtype=torch.int32, device=dev),
               csr_idx=torch.full((max(E, 1),), -7, dtype=torch.int32, device=dev))
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    rc = lib.dbaf_ba_tables_noalloc(p(ii), p(jj), E, n_frames, t0, t1, p(out["info"]), p(out["kx"]),
                                    p(out["kk_exp"]), p(out["csr_ptr"]), p(out["csr_idx"]),
                                    ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.dbaf_last_error()
    torch.cuda.synchronize()
    K = int(out["info"][0])
    n_valid = int(out["csr_ptr"][K])
    return dict(K=K, kx=out["kx"][:K].cpu().numpy(), kk_exp=out["kk_exp"].cpu().numpy(),
                csr_ptr=out["csr_ptr"][:K + 1].cpu().numpy(), csr_idx=out["csr_idx"][:n_valid].cpu().numpy(),
                info=out["info"].cpu().numpy())


def test_noalloc_tables_match_regular_entry_point_twice(dev, db):
    from droid_backends import _lib
    lib = _lib.lib
    rng = np.random.default_rng(42)
    # call 1: a large graph that fills much of the shared array; call 2: a small one right after
    for n_frames, n_edges in ((3000, 40000), (9, 17)):
        t1 = n_frames
        t0 = 1
        ii = torch.from_numpy(rng.integers(0, t1, n_edges)).to(dev)
        jj = torch.from_numpy(rng.integers(0, t1, n_edges)).to(dev)
        got = _noalloc_tables(lib, ii, jj, t0, t1, n_frames, dev)
        want = db.ba_tables(ii, jj, t0, t1, n_frames)
        assert got["K"] == int(want["info"][0]) and got["info"][2] == 0
        assert np.array_equal(got["kx"], want["kx"].numpy())
        assert np.array_equal(got["kk_exp"], want["kk_exp"].numpy())
        assert np.array_equal(got["csr_ptr"], want["csr_ptr"].numpy())
        assert np.array_equal(got["csr_idx"], want["csr_idx"].numpy())


def _numpy_tables(ii, jj, t0, t1):
    """The same tables from first principles: kx = sorted unique of cat(arange(t0,t1), ii) with its inverse map
    (src/droid_kernels.cu:1416-1424), and the edges grouped by source frame in ascending edge order."""
    ii = np.asarray(ii, np.int64)
    P = t1 - t0
    kx, inv = np.unique(np.concatenate([np.arange(t0, t1, dtype=np.int64), ii]), return_inverse=True)
    k_of_edge = inv[P:]
    order = np.argsort(k_of_edge, kind="stable")
    ptr = np.concatenate([[0], np.cumsum(np.bincount(k_of_edge, minlength=len(kx)))])
    return kx, inv, ptr, order


def test_noalloc_tables_match_numpy_at_the_size_limits(dev):
    """Exactly at the advertised limits (16 384 frames, 262 144 edges), where the carve-up of the fixed array reaches
    its largest offsets, and compared with numpy rather than with the other entry point (which runs the same
    kernel); then a small graph again to show that nothing of the large call survives in the shared array."""
    from droid_backends import _lib
    lib = _lib.lib
    rng = np.random.default_rng(7)
    for n_frames, n_edges, t0 in ((16384, 262144, 3), (12, 30, 2)):
        t1 = n_frames
        ii_h = rng.integers(0, n_frames, n_edges)
        jj_h = rng.integers(0, n_frames, n_edges)
        got = _noalloc_tables(lib, torch.from_numpy(ii_h).to(dev), torch.from_numpy(jj_h).to(dev), t0, t1, n_frames, dev)
        kx, inv, ptr, order = _numpy_tables(ii_h, jj_h, t0, t1)
        assert got["K"] == len(kx) and got["info"][2] == 0
        assert np.array_equal(got["kx"], kx)
        assert np.array_equal(got["kk_exp"], inv)
        assert np.array_equal(got["csr_ptr"], ptr)
        assert np.array_equal(got["csr_idx"], order)


def test_noalloc_rejects_sizes_beyond_its_fixed_array(dev):
    from droid_backends import _lib
    lib = _lib.lib
    z = torch.zeros(4, dtype=torch.int64, device=dev)
    o = torch.zeros(64, dtype=torch.int32, device=dev)
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    rc = lib.dbaf_ba_tables_noalloc(p(z), p(z), 4, 16385, 0, 2, p(o), p(o), p(o), p(o), p(o), None)
    assert rc == 1 and b"16384" in lib.dbaf_last_error()
    rc = lib.dbaf_ba_tables_noalloc(p(z), p(z), 262145, 8, 0, 2, p(o), p(o), p(o), p(o), p(o), None)
    assert rc == 1 and b"262144" in lib.dbaf_last_error()
