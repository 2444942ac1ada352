import os, sys, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "dba-fusion_b200")]
from dbaf_b200 import synth, global_ba
dev = torch.device("cuda:0")
cfg = sys.argv[1] if len(sys.argv) > 1 else "c5"
win = synth.make_window(cfg, seed=0, graph="banded", pose_noise=0.005)
t = {k: (torch.from_numpy(v.copy()).to(dev) if isinstance(v, np.ndarray) else v) for k, v in win.items()}
eng = global_ba.CudaEngine(t["poses"], t["disps"], t["intrinsics"], t["disps_sens"], t["targets"], t["weights"], t["eta"], t["ii"], t["jj"], win["t0"], win["t1"])
gba = global_ba.GlobalBA(eng, win["ii"], win["jj"], win["t0"], win["t1"])
for _ in range(int(sys.argv[2]) if len(sys.argv) > 2 else 2):
    gba.iterate()
torch.cuda.synchronize()
print("ok", eng.w.info().cpu()[:5].tolist())
