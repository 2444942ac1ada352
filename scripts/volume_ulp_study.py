"""How far is the reference's correlation volume (torch.matmul, fp16 in, fp16 out) from the exactly
rounded result, and how far is ours?  Prints fp16-ulp error histograms against a float64 product.
Run on the GPU box; the output backs the tolerance stated in tests/test_gpu_parity.py."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "dba-fusion_b200")]
import droid_backends as db  # noqa: E402
from dbaf_b200 import synth  # noqa: E402

dev = torch.device("cuda:0")
for cfg in ("tiny", "c2"):
    w = synth.make_window(cfg, seed=0)
    f = torch.from_numpy(synth.make_features(cfg, seed=0, n_frames=w["poses"].shape[0])).to(dev)
    f1, f2 = f[w["ii"][:2]].contiguous(), f[w["jj"][:2]].contiguous()
    n, C, ht, wd = f1.shape
    a = (f1.reshape(n, C, -1) / 4.0)
    b = (f2.reshape(n, C, -1) / 4.0)
    cublas = torch.matmul(a.transpose(1, 2), b)
    exact = torch.matmul(a.double().transpose(1, 2), b.double())
    ours = db.corr_volume_pyramid(f1, f2, 1)[0].reshape(n, ht * wd, ht * wd)
    rn = exact.half()
    # ulp of the correctly rounded value
    ulp = torch.pow(2.0, torch.floor(torch.log2(rn.double().abs().clamp_min(2.0 ** -14))) - 10)
    for name, x in (("cublas", cublas), ("ours", ours)):
        e = ((x.double() - exact).abs() / ulp)
        d = ((x.double() - rn.double()).abs() / ulp).round()
        print(cfg, name, "max err vs exact (ulp) %.3f" % e.max().item(),
              "hist of |x - rn(exact)| in ulp:", [int((d == k).sum()) for k in range(4)])
    dd = ((cublas.double() - ours.double()).abs() / ulp).round()
    print(cfg, "cublas vs ours ulp hist", [int((dd == k).sum()) for k in range(4)])
    for flag in (True, False):
        torch.backends.cuda.matmul.allow_fp16_reduced_precision_reduction = flag
        c2 = torch.matmul(a.transpose(1, 2), b)
        print(cfg, "allow_fp16_reduced_precision_reduction", flag, "max err (ulp) %.3f" % (((c2.double() - exact).abs() / ulp).max().item()))
