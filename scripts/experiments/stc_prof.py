import os, sys, numpy as np, torch
sys.path[:0] = ["/root/repo/dba-fusion_b200"]
from dbaf_b200 import synth, global_ba
dev = torch.device("cuda:0")
win = synth.make_window("c5", seed=0, graph="banded", pose_noise=0.005)
t = {k: (torch.from_numpy(v.copy()).to(dev) if isinstance(v, np.ndarray) else v) for k, v in win.items()}
eng = global_ba.CudaEngine(t["poses"], t["disps"], t["intrinsics"], t["disps_sens"], t["targets"], t["weights"], t["eta"], t["ii"], t["jj"], win["t0"], win["t1"])
for _ in range(2):
    eng.linearize(0.05, 0, -1)
torch.cuda.synchronize()
print("k_schur_tc transform warp 0, cycles per 32-pixel slice: wait raw, read raw, drain, wait operand stage, transform+store, fence+arrive:", eng.w.info().cpu()[10:16].tolist())
