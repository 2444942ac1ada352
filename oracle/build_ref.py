"""Builds the UNMODIFIED reference `droid_backends` CUDA extension for sm_100a into oracle/_ref/.

TEST INFRASTRUCTURE ONLY.  Sources are compiled where they lie (/root/reference/src, Eigen from
/root/reference/thirdparty/eigen); nothing is copied.  Only the module name differs
(-DTORCH_EXTENSION_NAME=droid_backends_ref) so that the reference and the B200 module can be
imported side by side, and oracle/ref_shim.h is force-included (see that file).

The result, oracle/_ref/droid_backends_ref.so, is git-ignored but travels to the GPU box with
the gpurun snapshot; /root/reference itself does not exist there.

usage: python oracle/build_ref.py [--force]
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("DBAF_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")
NAME = "droid_backends_ref"


def available() -> bool:
    return os.path.isdir(os.path.join(REF, "src"))


def build(force: bool = False) -> str | None:
    so = os.path.join(OUT, NAME + ".so")
    if os.path.exists(so) and not force:
        return so
    if not available():
        return None
    import torch
    T = os.path.dirname(torch.__file__)
    os.makedirs(os.path.join(OUT, "obj"), exist_ok=True)
    inc = [f"-I{T}/include", f"-I{T}/include/torch/csrc/api/include",
           f"-I{sysconfig.get_paths()['include']}", f"-I{REF}/thirdparty/eigen", f"-I{REF}/src",
           "-I/usr/local/cuda/include", f"-DTORCH_EXTENSION_NAME={NAME}", "-DTORCH_API_INCLUDE_EXTENSION_H",
           "-include", os.path.join(HERE, "ref_shim.h")]
    nvcc = ["/usr/local/cuda/bin/nvcc", "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-w",
            "-gencode=arch=compute_100a,code=sm_100a", "--expt-relaxed-constexpr"]
    cxx = ["g++", "-O3", "-std=c++17", "-fPIC", "-w"]
    jobs = []
    for src in ("droid_kernels.cu", "correlation_kernels.cu", "altcorr_kernel.cu"):
        obj = os.path.join(OUT, "obj", src + ".o")
        jobs.append((nvcc + inc + ["-c", os.path.join(REF, "src", src), "-o", obj], obj))
    obj = os.path.join(OUT, "obj", "droid.cpp.o")
    jobs.append((cxx + inc + ["-c", os.path.join(REF, "src", "droid.cpp"), "-o", obj], obj))

    def run(job):
        cmd, _ = job
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("reference build failed:\n" + " ".join(cmd) + "\n" + r.stderr[-4000:])

    with ThreadPoolExecutor(4) as ex:
        list(ex.map(run, jobs))
    link = ["g++", "-shared", "-o", so] + [o for _, o in jobs] + [
        f"-L{T}/lib", "-L/usr/local/cuda/lib64", "-lc10", "-ltorch", "-ltorch_cpu", "-ltorch_python",
        "-lc10_cuda", "-ltorch_cuda", "-lcudart", f"-Wl,-rpath,{T}/lib"]
    r = subprocess.run(link, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("reference link failed:\n" + r.stderr[-4000:])
    return so


def build_lietorch(force: bool = False) -> str | None:
    """The reference's lietorch extension (thirdparty/lietorch, unmodified, compiled where it lies) into
    oracle/_ref/lietorch_backends.so: what the reference's pure-PyTorch CPU path (dbaf/geom/ba.py) needs.  Used by
    scripts/time_geom_ba_cpu.py in the dev container only -- the Python half of that path lives under /root/reference
    and does not exist on the GPU box.  ~5 minutes (nvcc on lietorch_gpu.cu; the CPU path links against it)."""
    so = os.path.join(OUT, "lietorch_backends.so")
    if os.path.exists(so) and not force:
        return so
    lt = os.path.join(REF, "thirdparty", "lietorch")
    if not os.path.isdir(lt):
        return None
    import torch
    T = os.path.dirname(torch.__file__)
    os.makedirs(os.path.join(OUT, "obj"), exist_ok=True)
    inc = [f"-I{T}/include", f"-I{T}/include/torch/csrc/api/include", f"-I{sysconfig.get_paths()['include']}",
           f"-I{lt}/lietorch/include", f"-I{lt}/eigen", "-I/usr/local/cuda/include",
           "-DTORCH_EXTENSION_NAME=lietorch_backends", "-DTORCH_API_INCLUDE_EXTENSION_H",
           "-include", os.path.join(HERE, "ref_shim.h")]
    nvcc = ["/usr/local/cuda/bin/nvcc", "-O2", "-std=c++17", "-Xcompiler", "-fPIC", "-w",
            "-gencode=arch=compute_100a,code=sm_100a", "--expt-relaxed-constexpr"]
    cxx = ["g++", "-O2", "-std=c++17", "-fPIC", "-w"]
    jobs = []
    for src, tool in (("lietorch_gpu.cu", nvcc), ("lietorch_cpu.cpp", cxx), ("lietorch.cpp", cxx)):
        obj = os.path.join(OUT, "obj", src + ".o")
        jobs.append((tool + inc + ["-c", os.path.join(lt, "lietorch", "src", src), "-o", obj], obj))

    def run(job):
        r = subprocess.run(job[0], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("lietorch build failed:\n" + " ".join(job[0]) + "\n" + r.stderr[-4000:])

    with ThreadPoolExecutor(3) as ex:
        list(ex.map(run, jobs))
    link = ["g++", "-shared", "-o", so] + [o for _, o in jobs] + [
        f"-L{T}/lib", "-L/usr/local/cuda/lib64", "-lc10", "-ltorch", "-ltorch_cpu", "-ltorch_python",
        "-lc10_cuda", "-ltorch_cuda", "-lcudart", f"-Wl,-rpath,{T}/lib"]
    r = subprocess.run(link, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("lietorch link failed:\n" + r.stderr[-4000:])
    return so


def load():
    """Imports oracle/_ref/droid_backends_ref.so (None if it was never built)."""
    so = os.path.join(OUT, NAME + ".so")
    if not os.path.exists(so):
        return None
    import importlib.util
    import torch  # noqa: F401  (libtorch must be loaded first)
    spec = importlib.util.spec_from_file_location(NAME, so)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    path = build(force="--force" in sys.argv)
    print(path if path else "reference sources not available; nothing built")
