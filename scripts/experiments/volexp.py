import os, sys, torch, statistics
ROOT = "/root/repo"
sys.path[:0] = [os.path.join(ROOT, "dba-fusion_b200")]
import droid_backends as db
from dbaf_b200 import synth
w = synth.make_window("c2", seed=0)
f = torch.from_numpy(synth.make_features("c2", seed=0, n_frames=w["poses"].shape[0])).cuda()
ii, jj = torch.from_numpy(w["ii"]).cuda(), torch.from_numpy(w["jj"]).cuda()
buf = torch.empty(512 * 1024 * 1024 // 4, device="cuda")
ts = []
for it in range(8):
    buf.add_(1.0)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); db.corr_volume_pyramid_indexed(f, f, ii, jj, 4); b.record(); b.synchronize()
    if it >= 3: ts.append(a.elapsed_time(b))
print(sys.argv[1], "volume_pyramid_indexed ms", statistics.mean(ts))
