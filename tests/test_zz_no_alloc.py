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
