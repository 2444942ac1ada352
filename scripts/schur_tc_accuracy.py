"""Reduced system from the tensor-core Schur kernel vs the fp32 SIMT kernels, both against the float64 oracle."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "dba-fusion_b200"), os.path.join(ROOT, "oracle")]
import numpy as np, torch
import droid_backends as db
import dbaf_oracle as oracle
from dbaf_b200 import synth
dev = torch.device("cuda:0")
def relerr(a, b): return float(np.linalg.norm(np.asarray(a, np.float64) - b) / max(np.linalg.norm(b), 1e-300))
for cfg, reach in (("g40", 6), ("g24", 10)):
    w = synth.make_window(cfg, seed=3, graph="banded", reach=reach)
    out = {}
    for impl in ("tc", "simt"):
        db.set_option("schur_impl", "simt" if impl == "simt" else "auto")
        t = {k: (torch.from_numpy(v.copy()).to(dev) if isinstance(v, np.ndarray) else v) for k, v in w.items()}
        win = db._Window(t["poses"], t["disps"], t["intrinsics"], t["disps_sens"], t["targets"], t["weights"], t["eta"], t["ii"], t["jj"], w["t0"], w["t1"])
        win.prepare(); win.linearize(0.05)
        n = win.n; hv = win.Hv().cpu().numpy()
        H, v = hv[:n*n].reshape(n, n).copy(), hv[n*n:].copy()
        win.solve(1e-4, 0.1)
        out[impl] = (H, v, win.dx().cpu().numpy().reshape(-1).copy())
        win.linearize(0.05, motion_only=2)
        A = win.Hv().cpu().numpy()[:n*n].reshape(n, n).copy()
    tr = []
    p, d = w["poses"].astype(np.float64), w["disps"].astype(np.float64)
    oracle.ba(p, d, w["intrinsics"], w["disps_sens"], w["targets"], w["weights"], w["eta"], w["ii"], w["jj"], w["t0"], w["t1"], 1, 1e-4, 0.1, False, dtype=np.float64, trace=tr)
    H64, v64, dx64 = tr[0]["H"], tr[0]["v"], tr[0]["dx"].reshape(-1)
    S_tc, S_si = A - out["tc"][0], A - out["simt"][0]
    dg = np.arange(A.shape[0])
    print("norm A", np.linalg.norm(A), "norm S", np.linalg.norm(S_si), "norm H", np.linalg.norm(H64),
          "relerr S tc vs simt", relerr(S_tc, S_si), "diag rel", relerr(S_tc[dg, dg], S_si[dg, dg]),
          "mean signed diag", float(np.mean((S_tc[dg, dg] - S_si[dg, dg]) / S_si[dg, dg])))
    print(json.dumps({"cfg": cfg, "reach": reach, "HW": w["ht"] * w["wd"],
                      **{f"{k}_{i}": relerr(out[i][j], ref) for i in out for j, (k, ref) in enumerate((("H", H64), ("v", v64), ("dx", dx64)))}}))
