"""ctypes loader for libdbaf_b200.so (the C ABI declared in include/dbaf_b200.h).

There is no CPU or PyTorch fallback: if the library is missing the import fails loudly and
tells the user how to build it.
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.normpath(os.path.join(_HERE, "..", "csrc"))
LIB_PATH = os.path.join(CSRC, "libdbaf_b200.so")

c_int, c_float, c_void_p, c_size_t, c_int64 = (ctypes.c_int, ctypes.c_float, ctypes.c_void_p,
                                               ctypes.c_size_t, ctypes.c_int64)


class BaLayout(ctypes.Structure):
    """Mirror of `dbaf_ba_layout` (include/dbaf_b200.h)."""
    _fields_ = [(name, c_int64) for name in (
        "total_bytes", "info", "H", "v", "dx", "Lfac", "kx", "kk_exp", "csr_ptr", "csr_idx",
        "srow_ptr", "srow_idx", "frame2k", "cursor", "Q", "w", "Ei", "Eij", "fused", "exf", "rc_ptrs", "Kcap")]


# name -> (restype, argtypes); must list every symbol include/dbaf_b200.h declares
P = c_void_p
SIGNATURES = {
    "dbaf_version": (c_int, []),
    "dbaf_ba_layout_bytes": (c_int, []),
    "dbaf_last_error": (ctypes.c_char_p, []),
    "dbaf_set_option": (c_int, [ctypes.c_char_p, ctypes.c_char_p]),
    "dbaf_probe_fill": (c_int, [P, c_size_t, c_int, P]),
    "dbaf_proximity_base_count": (c_int, [c_int, c_int, c_int, c_int]),
    "dbaf_proximity_select": (c_int, [P, P, P, P, c_int, c_int, P, P, c_int, c_int, c_int, c_int, c_int, c_int, c_float,
                                      c_int, c_int, P, P, P, c_int, P, P]),
    "dbaf_filter_repeated_edges": (c_int, [P, P, P, c_int, P, P, c_int, P, P, P, P]),
    "dbaf_ba_schur": (c_int, [P, P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, P]),
    "dbaf_projmap": (c_int, [P, P, P, P, P, c_int, c_int, c_int, P, P, P]),
    "dbaf_frame_distance": (c_int, [P, P, P, P, P, c_int, c_int, c_int, c_float, P, P]),
    "dbaf_iproj": (c_int, [P, P, P, c_int, c_int, c_int, P, P]),
    "dbaf_depth_filter": (c_int, [P, P, P, P, P, c_int, c_int, c_int, c_int, P, P]),
    "dbaf_corr_index_forward": (c_int, [P, c_int, P, c_int, c_int, c_int, c_int, c_int, c_int, P, P]),
    "dbaf_corr_lookup_pyramid": (c_int, [ctypes.POINTER(c_void_p), c_int, c_int, P, c_int, c_int, c_int,
                                         c_int, c_int, c_int, P, P]),
    "dbaf_corr_index_backward": (c_int, [P, P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, P, P]),
    "dbaf_altcorr_forward": (c_int, [P, P, P] + [c_int] * 8 + [P, P]),
    "dbaf_altcorr_backward": (c_int, [P, P, P, P] + [c_int] * 8 + [P, P, P, P]),
    "dbaf_reproject": (c_int, [P, P, P, c_int, P, P, c_int, c_int, c_int, P, P, P]),
    "dbaf_corr_volume_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "dbaf_corr_volume_pyramid": (c_int, [P, P, c_int, c_int, c_int, c_int, c_int,
                                         ctypes.POINTER(c_void_p), P, c_size_t, P]),
    "dbaf_corr_volume_indexed_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int, c_int]),
    "dbaf_corr_volume_pyramid_indexed": (c_int, [P, c_int, P, c_int, P, P, c_int, c_int, c_int, c_int, c_int,
                                                 ctypes.POINTER(c_void_p), P, c_size_t, P]),
    "dbaf_corr_volume_status": (c_int, [P, c_size_t, P, P]),
    "dbaf_corr_volume_pyramid_pool": (c_int, [P, c_int, P, c_int, P, P, P, c_int, c_int, c_int, c_int, c_int, c_int,
                                              ctypes.POINTER(c_void_p), P, c_size_t, P]),
    "dbaf_corr_lookup_pyramid_pool": (c_int, [ctypes.POINTER(c_void_p), c_int, c_int, P, P] + [c_int] * 7 + [P, P]),
    "dbaf_ba_workspace_layout": (c_int, [c_int, c_int, c_int, c_int, c_int, c_int, ctypes.POINTER(BaLayout)]),
    "dbaf_ba_prepare": (c_int, [P, P, P, c_int, c_int, c_int, c_int, c_int, c_int, P]),
    "dbaf_ba_congruence": (c_int, [P, c_int, c_int, P, P, P, P]),
    "dbaf_ba_host_staging_bytes": (c_size_t, [c_int] * 7),
    "dbaf_ba_host": (c_int, [P, P, c_size_t] + [P] * 7 + [c_int, P, P] + [c_int] * 7 + [c_float, c_float, c_int, P, P, P]),
    "dbaf_ba_tables_noalloc": (c_int, [P, P, c_int, c_int, c_int, c_int, P, P, P, P, P, P]),
    "dbaf_ba_linearize": (c_int, [P, P, P, P, P, P, P, P, c_int, P, P, c_int, c_int, c_int, c_int, c_int,
                                  c_int, c_float, c_int, c_int, c_int, c_int, P]),
    "dbaf_ba_solve": (c_int, [P, c_int, c_int, c_float, c_float, P, c_int, c_int, P]),
    "dbaf_ba_solve_banded": (c_int, [P, c_int, c_int, c_float, c_float, c_int, P]),
    "dbaf_ba_solve_banded_launches": (c_int, [c_int, c_int]),
    "dbaf_ba_zero_system": (c_int, [P, c_int, c_int, c_int, P]),
    "dbaf_ba_band_strip_doubles": (c_size_t, [c_int, c_int]),
    "dbaf_ba_band_strip": (c_int, [P, c_int, c_int, c_int, P, c_int, P]),
    "dbaf_ba_retract": (c_int, [P, P, P, P, P, P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                c_int, c_int, P, P, P]),
    "dbaf_ba": (c_int, [P, P, P, P, P, P, P, P, c_int, P, P, c_int, c_int, c_int, c_int, c_int, c_int,
                        c_int, c_float, c_float, c_int, P, P, P]),
}


def load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"droid_backends (B200): {LIB_PATH} is missing. Build it with "
            f"`make -C {CSRC}` or `python -c 'import __graft_entry__ as g; g.build()'`. "
            "There is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)        # AttributeError if the library lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    if lib.dbaf_ba_layout_bytes() != ctypes.sizeof(BaLayout):
        raise ImportError(f"droid_backends (B200): {LIB_PATH} was built from a different include/dbaf_b200.h "
                          f"(dbaf_ba_layout is {lib.dbaf_ba_layout_bytes()} bytes there, {ctypes.sizeof(BaLayout)} here)")
    return lib


lib = load()


def check(rc: int, what: str):
    if rc != 0:
        msg = lib.dbaf_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")
