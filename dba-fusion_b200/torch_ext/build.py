"""Builds droid_backends_torch.so: the compiled torch-extension binding of the C ABI (host compiler only).

    python dba-fusion_b200/torch_ext/build.py [--force]

The result lies next to this file (git-ignored, travels with the gpurun snapshot) and links libdbaf_b200.so through
an $ORIGIN-relative rpath.  `import droid_backends_torch` then offers the reference module's eleven names."""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
NAME = "droid_backends_torch"


def build(force: bool = False) -> str:
    so = os.path.join(HERE, NAME + ".so")
    src = os.path.join(HERE, NAME + ".cpp")
    lib = os.path.join(ROOT, "dba-fusion_b200", "csrc", "libdbaf_b200.so")
    hdr = os.path.join(ROOT, "include", "dbaf_b200.h")
    if os.path.exists(so) and not force and os.path.getmtime(so) >= max(os.path.getmtime(src), os.path.getmtime(lib), os.path.getmtime(hdr)):
        return so
    import torch
    T = os.path.dirname(torch.__file__)
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-w", src, "-o", so,
           f"-I{T}/include", f"-I{T}/include/torch/csrc/api/include", f"-I{sysconfig.get_paths()['include']}",
           "-I/usr/local/cuda/include", f"-I{ROOT}/include", f"-DTORCH_EXTENSION_NAME={NAME}", "-DTORCH_API_INCLUDE_EXTENSION_H",
           f"-L{T}/lib", "-L/usr/local/cuda/lib64", f"-L{os.path.dirname(lib)}", "-ldbaf_b200", "-lc10", "-ltorch", "-ltorch_cpu",
           "-ltorch_python", "-lc10_cuda", "-ltorch_cuda", "-lcudart",
           f"-Wl,-rpath,{T}/lib", "-Wl,-rpath,$ORIGIN/../csrc"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("droid_backends_torch build failed:\n" + r.stderr[-4000:])
    return so


def load():
    """Imports droid_backends_torch.so (None if it was never built)."""
    so = os.path.join(HERE, NAME + ".so")
    if not os.path.exists(so):
        return None
    import importlib.util
    import torch  # noqa: F401  (libtorch must be loaded first)
    spec = importlib.util.spec_from_file_location(NAME, so)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
