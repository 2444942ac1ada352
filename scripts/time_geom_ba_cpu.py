"""CPU baseline of the reference itself: dbaf/geom/ba.py (`geom.BA`, pure PyTorch + lietorch) timed on this machine's
host cores on BASELINE.json configs[0] (8 keyframes, 24 edges, 64x64, 2 Gauss-Newton iterations, fixedp = 1).

DEV CONTAINER ONLY: it imports the reference's Python from /root/reference (which does not exist on the GPU box and
may not be copied into this repository) and the reference's lietorch extension built by oracle/build_ref.py
(build_lietorch) into oracle/_ref/.  Three stand-ins are needed to run that code on CPU tensors, none of them touching
its arithmetic: a `torch_scatter.scatter_sum` (index_add_ along dim 1), `device="cuda"` in the two stereo-baseline
literals of geom/projective_ops.py:105,136 mapped to the CPU, and lietorch found through sys.path.

    python scripts/time_geom_ba_cpu.py [--threads N]

Prints one JSON line (ms per iteration, threads) -- quoted in BASELINE.md next to the numpy port."""
import argparse
import json
import os
import statistics
import sys
import time
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("DBAF_REFERENCE", "/root/reference")
sys.path[:0] = [os.path.join(ROOT, "dba-fusion_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "oracle", "_ref"),
                os.path.join(REF, "thirdparty", "lietorch"), os.path.join(REF, "dbaf")]

import numpy as np  # noqa: E402
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--threads", type=int, default=os.cpu_count())
    ap.add_argument("--runs", type=int, default=20)
    args = ap.parse_args()
    torch.set_num_threads(args.threads)
    import build_ref
    if build_ref.build_lietorch() is None:
        raise SystemExit("the reference's lietorch sources are not available here")

    ts = types.ModuleType("torch_scatter")

    def scatter_sum(src, index, dim=1, dim_size=None):
        out = torch.zeros(*src.shape[:dim], dim_size, *src.shape[dim + 1:], dtype=src.dtype)
        return out.index_add_(dim, index, src)
    ts.scatter_sum = scatter_sum
    sys.modules["torch_scatter"] = ts
    as_tensor = torch.as_tensor
    torch.as_tensor = lambda *a, **k: as_tensor(*a, **{**k, "device": "cpu"} if k.get("device") == "cuda" else k)

    import lietorch
    from geom.ba import BA
    from dbaf_b200 import synth

    w = synth.make_window("c1", seed=0)
    P, ht, wd = w["poses"].shape[0], w["ht"], w["wd"]
    poses = lietorch.SE3(torch.from_numpy(w["poses"])[None])
    disps = torch.from_numpy(w["disps"])[None]
    intr = torch.from_numpy(w["intrinsics"])[None, None].repeat(1, P, 1)
    ii, jj = torch.from_numpy(w["ii"]), torch.from_numpy(w["jj"])
    target = torch.from_numpy(w["targets"])[None].permute(0, 1, 3, 4, 2).contiguous()      # [1,E,ht,wd,2]
    weight = torch.from_numpy(w["weights"])[None].permute(0, 1, 3, 4, 2).contiguous()
    kx = torch.unique(ii)
    eta = torch.from_numpy(w["eta"])[: len(kx)][None] if w["eta"].shape[0] >= len(kx) else torch.from_numpy(w["eta"])[None]

    def run():
        p, d = poses, disps
        for _ in range(2):
            p, d = BA(target, weight, eta, p, d, intr, ii, jj, fixedp=w["t0"])
        return p, d

    for _ in range(3):
        run()
    times = []
    for _ in range(args.runs):
        a = time.perf_counter()
        run()
        times.append((time.perf_counter() - a) * 1e3)
    print(json.dumps(dict(what="reference geom.BA (dbaf/geom/ba.py) on CPU, c1: 8 keyframes, 24 edges, 64x64, 2 iterations",
                          ms_per_iter=statistics.median(times) / 2, threads=torch.get_num_threads(), cores=os.cpu_count(),
                          runs=len(times))))


if __name__ == "__main__":
    main()
