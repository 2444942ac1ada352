"""Shared helpers for the parity tests."""
import numpy as np


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
