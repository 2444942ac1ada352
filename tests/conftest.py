import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "dba-fusion_b200"), os.path.join(ROOT, "oracle"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    # keep the no-allocation entry point's test (if present) last, everything else in file order
    items.sort(key=lambda it: (os.path.basename(str(it.fspath)).startswith("test_zz_"),))


@pytest.fixture(scope="session")
def oracle():
    import dbaf_oracle
    return dbaf_oracle


@pytest.fixture(scope="session")
def synth():
    from dbaf_b200 import synth as s
    return s


@pytest.fixture(scope="session")
def dev():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


@pytest.fixture(scope="session")
def db():
    import droid_backends
    return droid_backends


@pytest.fixture(autouse=True)
def _default_kernel_options():
    """Kernel-selection switches (droid_backends.set_option) never leak from one test into the next."""
    yield
    mod = sys.modules.get("droid_backends")
    if mod is not None:
        for key in ("schur_impl", "band_impl", "band_sweep", "volume_impl", "eij"):
            mod.set_option(key, "auto")
        mod.set_option("band_domains", "0")
        mod.set_strict(False)


@pytest.fixture(scope="session")
def dbt():
    """droid_backends as a compiled torch extension over the same C ABI (dba-fusion_b200/torch_ext)."""
    import importlib.util
    path = os.path.join(ROOT, "dba-fusion_b200", "torch_ext", "build.py")
    spec = importlib.util.spec_from_file_location("dbaf_torch_ext_build", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    ext = mod.load()
    if ext is None:
        pytest.skip("droid_backends_torch.so not built (python dba-fusion_b200/torch_ext/build.py)")
    return ext


@pytest.fixture(scope="session")
def ref():
    """The reference CUDA extension rebuilt for sm_100a (oracle/_ref/droid_backends_ref.so)."""
    import build_ref
    mod = build_ref.load()
    if mod is None:
        pytest.skip("oracle/_ref/droid_backends_ref.so not built (python oracle/build_ref.py)")
    return mod


def relerr(a, b):
    """normwise relative error max|a-b| / max|b|."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))
