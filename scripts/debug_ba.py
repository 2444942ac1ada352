import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "dba-fusion_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import droid_backends as db, dbaf_oracle as O
from dbaf_b200 import synth
from helpers import relerr, to_torch, ba_args
dev = torch.device("cuda:0")
cfg, kw = sys.argv[1], eval(sys.argv[2]) if len(sys.argv) > 2 else {}
w = synth.make_window(cfg, **kw)
for iters in (1, 2):
    for rep in range(2):
        t = to_torch(w, dev)
        dx, dz = db.ba(*ba_args(t), w["t0"], w["t1"], iters, 1e-4, 0.1, False)
        p64, d64 = w["poses"].astype(np.float64), w["disps"].astype(np.float64)
        tr = []
        dx64, dz64 = O.ba(p64, d64, w["intrinsics"], w["disps_sens"], w["targets"], w["weights"], w["eta"], w["ii"], w["jj"], w["t0"], w["t1"], iters, 1e-4, 0.1, False, dtype=np.float64, trace=tr)
        p32, d32 = w["poses"].copy(), w["disps"].copy()
        dx32, dz32 = O.ba(p32, d32, w["intrinsics"], w["disps_sens"], w["targets"], w["weights"], w["eta"], w["ii"], w["jj"], w["t0"], w["t1"], iters, 1e-4, 0.1, False, dtype=np.float32)
        print(cfg, "iters", iters, "rep", rep, "dx", relerr(dx.cpu().numpy(), dx64), "(f32 oracle", relerr(dx32, dx64), ") dz", relerr(dz.cpu().numpy(), dz64), "(", relerr(dz32, dz64), ") poses", relerr(t["poses"].cpu().numpy(), p64), "(", relerr(p32, p64), ") disps", relerr(t["disps"].cpu().numpy(), d64), "info", db.ba_info()[:5].tolist(),
              "cond", np.linalg.cond(tr[-1]["H"]))
