"""One fused volume + pyramid build at the C2 shape (for ncu captures)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "dba-fusion_b200")]
import torch
import droid_backends as db
from dbaf_b200 import synth
cfg = sys.argv[1] if len(sys.argv) > 1 else "c2"
w = synth.make_window(cfg, seed=0)
f = torch.from_numpy(synth.make_features(cfg, seed=0, n_frames=w["poses"].shape[0])).cuda()
f1, f2 = f[w["ii"]].contiguous(), f[w["jj"]].contiguous()
coords = torch.from_numpy(synth.lookup_coords(w, 0.05, 0)).cuda()
for _ in range(3):
    pyr = db.corr_volume_pyramid(f1, f2, 4)
    out = db.corr_lookup_pyramid(pyr, coords, 3)
torch.cuda.synchronize()
print("ok", out.shape)
