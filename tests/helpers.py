"""Shared helpers for the parity tests."""
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def record(test, **values):
    """Append the ACHIEVED error figures of a parity test to gpurun_out/parity_achieved.jsonl (scratch; the
    table in DESIGN.md section 4 is filled from it)."""
    out = os.path.join(ROOT, "gpurun_out")
    try:
        os.makedirs(out, exist_ok=True)
        with open(os.path.join(out, "parity_achieved.jsonl"), "a") as f:
            f.write(json.dumps(dict(test=test, **{k: (float(v) if v is not None else None) for k, v in values.items()})) + "\n")
    except OSError:
        pass


def relerr(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    if a.size == 0:
        return 0.0
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def to_torch(win, dev):
    import torch
    out = {}
    for k, v in win.items():
        if isinstance(v, np.ndarray):
            out[k] = torch.from_numpy(v.copy()).to(dev)
        else:
            out[k] = v
    return out


def ba_args(t):
    return (t["poses"], t["disps"], t["intrinsics"], t["disps_sens"], t["targets"], t["weights"],
            t["eta"], t["ii"], t["jj"])
