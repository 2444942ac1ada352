"""Where do the fused (k_linschur_tc) and the stored-Eij paths differ at the C5 size?  (diagnostic)"""
import os, sys, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "dba-fusion_b200")]
import droid_backends as db
from dbaf_b200 import synth
dev = torch.device("cuda:0")
cfg = sys.argv[1] if len(sys.argv) > 1 else "c5"
win = synth.make_window(cfg, seed=0, graph="banded", pose_noise=0.005)
def system(mode):
    db.set_option("eij", mode)
    t = {k: (torch.from_numpy(v.copy()).to(dev) if isinstance(v, np.ndarray) else v) for k, v in win.items()}
    w = db._Window(t["poses"], t["disps"], t["intrinsics"], t["disps_sens"], t["targets"], t["weights"], t["eta"], t["ii"], t["jj"], win["t0"], win["t1"])
    w.prepare(); w.linearize(0.05)
    n = w.n
    hv = w.Hv().clone()
    K = int(w.info().cpu()[0])
    Q = w._view(w.L.Q, K * w.HW, torch.float32).view(K, w.HW).clone()
    wv = w._view(w.L.w, K * w.HW, torch.float32).view(K, w.HW).clone()
    return hv[:n * n].view(n, n), hv[n * n:], Q, wv
Ha, va, Qa, wa = system("store")
Hb, vb, Qb, wb = system("recompute")
Hc, vc, Qc, wc = system("recompute")
rel = lambda a, b: float((a - b).abs().max() / b.abs().max())
print("H fused vs stored", rel(Hb, Ha), "v", rel(vb, va), "Q", rel(Qb, Qa), "w", rel(wb, wa))
print("H fused run-to-run", rel(Hc, Hb), "v", rel(vc, vb))
P = Ha.shape[0] // 6
D = (Hb - Ha).abs().view(P, 6, P, 6).amax(dim=(1, 3))
scale = float(Ha.abs().max())
idx = torch.nonzero(D > 1e-5 * scale)
print("blocks above 1e-5:", idx.shape[0], "of", int((Ha.view(P, 6, P, 6).abs().amax(dim=(1, 3)) > 0).sum()))
print("first:", idx[:20].tolist())
print("diag-block errs:", float(torch.diagonal(D).max() / scale), " offdiag:", float((D - torch.diag(torch.diagonal(D))).max() / scale))
dq = ((Qb - Qa).abs() / Qa.abs().clamp(min=1e-30)).amax(dim=1)
print("frames with Q rel err > 1e-4:", torch.nonzero(dq > 1e-4).flatten()[:20].tolist(), float(dq.max()))
dw = (wb - wa).abs().amax(dim=1) / wa.abs().max()
print("frames with w err > 1e-5:", torch.nonzero(dw > 1e-5).flatten()[:20].tolist(), float(dw.max()))
