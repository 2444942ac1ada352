"""One pass of the whole hot path at the C2 shape (for ncu captures): ba x2 iterations, indexed volume + pyramid,
4-level lookup, altcorr, projmap / frame_distance."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "dba-fusion_b200")]
import numpy as np, torch
import droid_backends as db
from dbaf_b200 import synth
dev = torch.device("cuda:0")
cfg = sys.argv[1] if len(sys.argv) > 1 else "c2"
w = synth.make_window(cfg, seed=0)
t = {k: (torch.from_numpy(v.copy()).to(dev) if isinstance(v, np.ndarray) else v) for k, v in w.items()}
f = torch.from_numpy(synth.make_features(cfg, seed=0, n_frames=w["poses"].shape[0])).to(dev)
coords = torch.from_numpy(synth.lookup_coords(w, 0.05, 0)).to(dev)
for _ in range(int(sys.argv[2]) if len(sys.argv) > 2 else 2):
    p, d = t["poses"].clone(), t["disps"].clone()
    db.ba(p, d, t["intrinsics"], t["disps_sens"], t["targets"], t["weights"], t["eta"], t["ii"], t["jj"],
          w["t0"], w["t1"], 2, 1e-4, 0.1, False)
    pyr = db.corr_volume_pyramid_indexed(f, f, t["ii"], t["jj"], 4)
    out = db.corr_lookup_pyramid(pyr, coords, 3)
    db.frame_distance(t["poses"], t["disps"], t["intrinsics"], t["ii"], t["jj"], 0.3)
torch.cuda.synchronize()
print("ok", out.shape)
