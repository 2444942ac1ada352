"""The C-ABI library loads on a machine without a GPU, exports every symbol that
include/dbaf_b200.h declares, and its host-only entry points validate arguments."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "dbaf_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(dbaf_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(db):
    from droid_backends import _lib
    names = _declared()
    assert len(names) >= 19
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), f"{n} is declared in include/dbaf_b200.h but not exported"
    assert set(names) == set(_lib.SIGNATURES), "ctypes signature table out of sync with the header"
    assert _lib.lib.dbaf_version() >= 100


def test_module_surface_matches_reference(db):
    # the 11 names registered at src/droid.cpp:297-316
    for name in ("ba", "ba_extend", "frame_distance", "projmap", "depth_filter", "iproj",
                 "altcorr_forward", "altcorr_backward", "corr_index_forward", "corr_index_backward", "BACore"):
        assert hasattr(db, name)
    core = db.BACore()
    for m in ("init", "hessian", "optimize", "retract"):
        assert callable(getattr(core, m))


def test_layout_and_argument_errors(db):
    from droid_backends import _lib
    L = _lib.BaLayout()
    assert _lib.lib.dbaf_ba_workspace_layout(16, 40, 1, 12, 64, 64, ctypes.byref(L)) == 0
    n = 6 * 11
    assert L.v == L.H + n * n * 8              # H and v are one contiguous span
    assert L.Kcap == 16 and L.total_bytes > L.Eij + 40 * 6 * 4096 * 4 - 1
    offs = [L.info, L.H, L.dx, L.Lfac, L.kx, L.kk_exp, L.csr_ptr, L.csr_idx, L.srow_ptr, L.srow_idx,
            L.frame2k, L.cursor, L.Q, L.w, L.Ei, L.Eij]
    assert offs == sorted(offs) and all(o % 256 == 0 for o in offs)
    # bad arguments are rejected with a message, nothing is launched
    assert _lib.lib.dbaf_ba_workspace_layout(16, 40, 5, 3, 64, 64, ctypes.byref(L)) == 1
    assert b"t0" in _lib.lib.dbaf_last_error()
    assert _lib.lib.dbaf_corr_index_forward(None, 7, None, 1, 4, 4, 4, 4, 3, None, None) == 1
    assert b"dtype" in _lib.lib.dbaf_last_error()
    assert _lib.lib.dbaf_corr_volume_pyramid(None, None, 1, 128, 8, 8, 9, None, None, 0, None) == 1


def test_contiguity_errors_like_the_reference(db):
    import pytest
    import torch
    t = torch.zeros(4, 7)
    with pytest.raises(RuntimeError, match="must be contiguous"):
        db.frame_distance(t.t(), t, t, t.long(), t.long(), 0.3)
    with pytest.raises(RuntimeError, match="CUDA"):
        db.iproj(t, torch.zeros(4, 2, 2), torch.zeros(4))


def test_compiled_torch_extension_exports_the_reference_surface(dbt):
    """The torch-extension build of the binding (what src/droid.cpp becomes) loads without a GPU and registers the
    eleven names of src/droid.cpp:297-316."""
    for name in ("ba", "ba_extend", "frame_distance", "projmap", "depth_filter", "iproj",
                 "altcorr_forward", "altcorr_backward", "corr_index_forward", "corr_index_backward", "BACore"):
        assert hasattr(dbt, name)
    core = dbt.BACore()
    for m in ("init", "hessian", "optimize", "retract"):
        assert callable(getattr(core, m))
    import pytest
    import torch
    t = torch.zeros(4, 7)
    with pytest.raises(RuntimeError, match="must be contiguous"):
        dbt.frame_distance(t.t(), t, t, t.long(), t.long(), 0.3)
