"""Golden vectors for `reproject` from the reference's own Python (dbaf/geom/projective_ops.py), run on CPU in the
development container.  `lietorch` (a compiled extension that is not installed) is replaced by the small SE3 class
below, which restates lietorch's quaternion algebra (thirdparty/lietorch/lietorch/include/se3.h: inv, operator*,
act4); everything else -- iproj, proj, the stereo special case, the validity test, the tensor indexing -- is the
reference's code, unmodified.  Its one hard-coded device="cuda" literal (projective_ops.py:105) is redirected to the
CPU by wrapping torch.as_tensor for the duration of the call.

usage: python tests/golden/make_golden_reproject.py   (needs /root/reference; writes tests/golden/ref_reproject.npz)
"""
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "dba-fusion_b200")]


def _qmul(a, b):
    ax, ay, az, aw = a.unbind(-1)
    bx, by, bz, bw = b.unbind(-1)
    return torch.stack([aw * bx + ax * bw + ay * bz - az * by, aw * by - ax * bz + ay * bw + az * bx,
                        aw * bz + ax * by - ay * bx + az * bw, aw * bw - ax * bx - ay * by - az * bz], -1)


def _qrot(q, v):
    qv = q[..., :3]
    uv = 2.0 * torch.cross(qv, v, dim=-1)
    return v + q[..., 3:4] * uv + torch.cross(qv, uv, dim=-1)


class SE3:
    def __init__(self, data):
        self.data = data

    def inv(self):
        q = self.data[..., 3:] * torch.tensor([-1.0, -1.0, -1.0, 1.0])
        return SE3(torch.cat([-_qrot(q, self.data[..., :3]), q], -1))

    def __mul__(self, other):
        if isinstance(other, SE3):
            q = _qmul(self.data[..., 3:], other.data[..., 3:])
            t = self.data[..., :3] + _qrot(self.data[..., 3:], other.data[..., :3])
            return SE3(torch.cat([t, q], -1))
        X = other                                    # homogeneous points [..., 4]
        Y = _qrot(self.data[..., 3:], X[..., :3]) + self.data[..., :3] * X[..., 3:4]
        return torch.cat([Y, X[..., 3:4]], -1)

    def __getitem__(self, idx):
        return SE3(self.data[idx])


def main():
    stub = types.ModuleType("lietorch")
    stub.SE3, stub.Sim3 = SE3, type("Sim3", (), {})
    sys.modules["lietorch"] = stub
    sys.path.insert(0, "/root/reference/dbaf")
    from geom import projective_ops as pops
    from dbaf_b200 import synth
    real_as_tensor = torch.as_tensor
    torch.as_tensor = lambda data, **kw: real_as_tensor(data, **{k: v for k, v in kw.items() if k != "device"})
    out = {}
    try:
        for name, cfg, kw in (("tiny", "tiny", dict(stereo_edges=2, near_frac=0.05)), ("c3", "c3", dict(stereo_edges=1))):
            w = synth.make_window(cfg, seed=11, **kw)
            N = w["poses"].shape[0]
            rng = np.random.default_rng(5)
            K = np.tile(w["intrinsics"], (N, 1)) * rng.uniform(0.9, 1.1, (N, 4)).astype(np.float32)
            sel = slice(-12, None)                    # the stereo edges are appended at the end
            ii, jj = w["ii"][sel], w["jj"][sel]
            coords, valid = pops.projective_transform(SE3(torch.from_numpy(w["poses"])[None]),
                                                      torch.from_numpy(w["disps"])[None], torch.from_numpy(K)[None],
                                                      torch.from_numpy(ii), torch.from_numpy(jj))
            for k, v in (("poses", w["poses"]), ("disps", w["disps"]), ("intr", K), ("ii", ii), ("jj", jj),
                         ("coords", coords[0].numpy()), ("valid", valid[0].numpy())):
                out[f"{name}_{k}"] = v
    finally:
        torch.as_tensor = real_as_tensor
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "ref_reproject.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
