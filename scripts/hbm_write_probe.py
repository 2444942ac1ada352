"""HBM write ceilings on this box: cudaMemsetAsync, a linear 128-bit-store kernel, and the tiled store patterns of the
correlation-volume kernel (128 rows of 8 KB written in 512 B .. 8 KB pieces per row).  python scripts/hbm_write_probe.py"""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "dba-fusion_b200")]
import droid_backends as db
x = torch.empty(1 << 31, dtype=torch.uint8, device="cuda")
names = {0: "cudaMemsetAsync", 1: "linear st.v4", 2: "tiled 512 B", 3: "tiled 1 KB", 4: "tiled 2 KB", 5: "tiled 8 KB"}
for kind in range(6):
    best = 1e9
    for _ in range(7):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); db.probe_fill(x, kind); b.record(); b.synchronize()
        best = min(best, a.elapsed_time(b))
    print(f"{names[kind]:18s} {x.numel() / best / 1e6:8.1f} GB/s")
