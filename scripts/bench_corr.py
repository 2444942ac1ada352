"""Correlation half of the metric on one GPU: volume + pyramid build (ours vs the reference's torch.matmul +
3x avg_pool2d, dbaf/modules/corr.py:24-38,63-71) and the 4-level lookup (ours vs reference kernel when
oracle/_ref is present).  Prints one JSON line."""
import json, os, sys, statistics
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "dba-fusion_b200"), os.path.join(ROOT, "oracle")]
import droid_backends as db
from dbaf_b200 import synth

def timeit(fn, steps=10, warm=3, flush=None):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    out = []
    for _ in range(steps):
        if flush is not None: flush.add_(1.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize()
        out.append(a.elapsed_time(b))
    return statistics.mean(out)

cfg = sys.argv[1] if len(sys.argv) > 1 else "c2"
dev = torch.device("cuda:0")
w = synth.make_window(cfg, seed=0)
f = torch.from_numpy(synth.make_features(cfg, seed=0, n_frames=w["poses"].shape[0])).to(dev)
f1, f2 = f[w["ii"]].contiguous(), f[w["jj"]].contiguous()
coords = torch.from_numpy(synth.lookup_coords(w, 0.05, 0)).to(dev)
flush = torch.empty(128 * 1024 * 1024, device=dev)
n, C, ht, wd = f1.shape
def ref_volume():
    a = f1.reshape(n, C, ht * wd) / 4.0
    b = f2.reshape(n, C, ht * wd) / 4.0
    corr = torch.matmul(a.transpose(1, 2), b).reshape(n * ht * wd, 1, ht, wd)
    pyr = []
    for i in range(4):
        pyr.append(corr.view(n, ht, wd, ht // 2 ** i, wd // 2 ** i))
        corr = torch.nn.functional.avg_pool2d(corr, 2, stride=2)
    return pyr
res = {"config": cfg, "edges": n}
res["volume_ms_ours"] = timeit(lambda: db.corr_volume_pyramid(f1, f2, 4), flush=flush)
res["volume_ms_ref_torch"] = timeit(ref_volume, flush=flush)
# the call the pipeline makes (dbaf/covisible_graph.py:129-131): gather by edge, then CorrBlock
ii_t, jj_t = torch.from_numpy(w["ii"]).to(dev), torch.from_numpy(w["jj"]).to(dev)
res["volume_indexed_ms_ours"] = timeit(lambda: db.corr_volume_pyramid_indexed(f, f, ii_t, jj_t, 4), flush=flush)
def ref_gather_volume():
    global f1, f2
    f1, f2 = f[ii_t], f[jj_t]
    return ref_volume()
res["volume_indexed_ms_ref_torch"] = timeit(ref_gather_volume, flush=flush)
pyr = db.corr_volume_pyramid(f1, f2, 4)
res["lookup_ms_ours_fused"] = timeit(lambda: db.corr_lookup_pyramid(pyr, coords, 3), flush=flush)
def per_level(mod):
    return [mod.corr_index_forward(pyr[l], (coords / 2 ** l).contiguous(), 3)[0] for l in range(4)]
res["lookup_ms_ours_per_level"] = timeit(lambda: per_level(db), flush=flush)
try:
    import build_ref
    ref = build_ref.load()
    if ref is not None:
        res["lookup_ms_ref"] = timeit(lambda: per_level(ref), flush=flush)
except Exception as ex:
    res["ref_error"] = str(ex)
print(json.dumps(res))
