"""Write-only / copy HBM bandwidth on this box (context for the volume kernel's roofline)."""
import torch, json
def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize()
        best = min(best, a.elapsed_time(b))
    return best
res = {}
for gb in (2, 8):
    x = torch.empty(gb * (1 << 30) // 2, dtype=torch.float16, device="cuda")
    y = torch.empty_like(x)
    res[f"fill_{gb}GB_GBps"] = gb * (1 << 30) / t(lambda: x.fill_(1.0)) / 1e6
    res[f"memset_{gb}GB_GBps"] = gb * (1 << 30) / t(lambda: x.zero_()) / 1e6
    res[f"copy_{gb}GB_GBps_rw"] = 2 * gb * (1 << 30) / t(lambda: y.copy_(x)) / 1e6
    res[f"sum_{gb}GB_GBps_read"] = gb * (1 << 30) / t(lambda: x.view(torch.int16).sum()) / 1e6
    del x, y
print(json.dumps(res))
