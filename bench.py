#!/usr/bin/env python
"""Benchmark of the dense-BA / correlation hot path (BASELINE.json: "dense-BA ms/iter + corr edges/s vs ref CUDA &
CPU; 1/2/4/8 B200 global-BA scaling").

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload c2|c2_real|c3|c4|c5]

The line of record is the SAME workload at every N, so that the 1/2/4/8 series is one series: global / backend BA
over a 1000-keyframe banded graph (BASELINE.json configs[4], the largest single-GPU configuration and the only
one that shards).  One step = one Gauss-Newton iteration; `value` is milliseconds per iteration with all inputs
resident in HBM (CUDA events, max over ranks, L2 flushed between steps), `e2e` the same iteration from / to pinned
HOST memory through dbaf_b200.global_ba.iterate_host.  For N > 1 the depth frames are sharded over the ranks
(strong scaling, NCCL).

At N = 1 the line also carries the frontend measurements as named sub-records: `frontend_c2` (BASELINE.json
configs[1]: 64x64, 12 keyframes, 48 edges, `droid_backends.ba` with 2 iterations, device-resident and end to end
through `ba_host`, the correlation volume / lookup / reproject kernels, the unmodified reference on the same GPU),
`frontend_c2_real` (55x55, what the TUM-VI demo really produces), `frontend_c3` (28x107, KITTI-360), `bacore` (the
IMU-mode call sequence BACore.init -> 2 x (hessian -> host -> retract) and the marginalisation call), `c4_altcorr`
(on-the-fly correlation against the materialised volume), and `cpu_baseline`.

--impl reference runs the unmodified reference extension (oracle/_ref/droid_backends_ref.so, built by
oracle/build_ref.py from the reference's own sources) on the same workload; if that library is absent it times the
oracle's CPU port instead.  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "dba-fusion_b200"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402
import torch  # noqa: E402


# ----------------------------------------------------------------------------- helpers
# DRAM bytes per launch read from committed ncu captures (profiles/), keyed by (workload, kernel)
NCU_TRAFFIC_BYTES = {("c2", "k_linearize2"): 3.79e6, ("c2", "k_volume_tc"): 2.116e9, ("c2", "k_corr_lookup"): 4.14e8}   # profiles/r2_ncu_c2.md


def write_only_peak(dev):
    """HBM write-only bandwidth measured here and now (GB/s): the best of cudaMemsetAsync and the library's
    hand-written 128-bit-store fill over 2 GiB (dbaf_probe_fill), best of 7 each.  Reported next to -- never instead
    of -- the fraction of the copy peak in MEASURED_PEAKS.json."""
    import droid_backends as db
    x = torch.empty(1 << 31, dtype=torch.uint8, device=dev)
    out = {}
    for kind in (0, 1):
        best = 1e9
        for _ in range(7):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            db.probe_fill(x, kind)
            b.record()
            b.synchronize()
            best = min(best, a.elapsed_time(b))
        out["memset" if kind == 0 else "st128"] = x.numel() / (best * 1e-3) / 1e9
    return max(out.values()), out


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return dict(hbm=float(d["hbm_gbs"]), tf_burst=float(d["bf16_tflops"]),
                    tf_sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, src="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return dict(sm_mhz=statistics.median(sm) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=reasons, samples=len(sm))


class L2Flusher:
    def __init__(self, dev, mb=512):
        self.buf = torch.empty(mb * 1024 * 1024 // 4, dtype=torch.float32, device=dev)

    def __call__(self):
        self.buf.add_(1.0)          # read + write of a buffer 4x the 126 MB L2


def timed_steps(fn, steps, warmup, flush, reset=None):
    """Device time of `steps` calls of fn (CUDA events on the current stream), L2 flushed and
    state reset between steps outside the timed intervals.  Returns per-step milliseconds."""
    for _ in range(warmup):
        if reset:
            reset()
        fn()
    torch.cuda.synchronize()
    out = []
    for _ in range(steps):
        if reset:
            reset()
        flush()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        out.append(a.elapsed_time(b))
    return out


def window_tensors(win, dev):
    return {k: (torch.from_numpy(v.copy()).to(dev) if isinstance(v, np.ndarray) else v) for k, v in win.items()}


def ba_call_args(t):
    return (t["poses"], t["disps"], t["intrinsics"], t["disps_sens"], t["targets"], t["weights"], t["eta"],
            t["ii"], t["jj"])


def global_ba_config():
    from dbaf_b200 import global_ba
    return global_ba.c5_config()


def frontend_config(cfg, n_frames, n_edges, ht, wd, iters):
    """`config` of a frontend line; both arms print the identical dict."""
    return dict(workload=f"{cfg}: {ht}x{wd} disparity maps, {n_frames} keyframes, {n_edges} edges, "
                         f"{iters} GN iterations per step (droid_backends.ba)",
                l2="512 MB read+write between steps (inputs are smaller than L2)")


# --------------------------------------------------------------------- single-GPU arm
def bench_local(args, dev, rank, world):
    import droid_backends as db
    from dbaf_b200 import synth
    cfg = args.workload
    ht, wd, P, E, _ = synth.CONFIGS[cfg]
    win = synth.make_window(cfg, seed=0)
    t = window_tensors(win, dev)
    p0, d0 = t["poses"].clone(), t["disps"].clone()
    t0, t1, iters = win["t0"], win["t1"], 2
    flush = L2Flusher(dev)
    peaks = measured_peaks()
    HW, K = ht * wd, win["eta"].shape[0]

    def reset():
        t["poses"].copy_(p0)
        t["disps"].copy_(d0)

    def step():
        db.ba(*ba_call_args(t), t0, t1, iters, 1e-4, 0.1, False)

    with ClockSampler(dev.index or 0) as clk:
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()
        ms = timed_steps(step, args.steps, args.warmup, flush, reset)
        torch.cuda.synchronize()
        if world > 1:
            torch.distributed.barrier()
    step_ms = sum(ms) / len(ms)

    # ---- per-stage device times (same inputs, L2 flushed): which kernel dominates
    w = db._Window(*ba_call_args(t), t0, t1)
    w.prepare()
    stage = {}
    for name, fn in (("linearize", lambda: w.linearize(0.05, motion_only=2, zero_first=False)),
                     ("linearize+schur", lambda: w.linearize(0.05)),
                     ("solve", lambda: w.solve(1e-4, 0.1)),
                     ("retract", lambda: w.retract(want_dx=False, want_dz=False))):
        stage[name] = statistics.mean(timed_steps(fn, max(5, args.steps), 3, flush, reset))
    lin_ms = stage["linearize"]
    lin_bytes = E * HW * 20 + K * HW * 12          # algorithmic bytes of the linearise kernel
    # traffic: dram__bytes_read.sum + dram__bytes_write.sum of one k_linearize2 launch on this workload from the
    # committed `ncu --set full` capture (profiles/r2_ncu_c2.md); it cannot be measured outside a profiler
    traffic = NCU_TRAFFIC_BYTES.get((cfg, "k_linearize2"))
    roof = dict(bound="hbm", kernel="k_linearize2", achieved=lin_bytes / (lin_ms * 1e-3) / 1e9, peak=peaks["hbm"],
                unit="GB/s", peak_source=peaks["src"] + " (burst, kernel timed alone)", traffic=traffic,
                algorithmic_bytes=lin_bytes, ms=lin_ms,
                note="latency-bound at this size: 48 edges x 4096 px is 4.5 MB, far below what fills the machine")
    roof["frac"] = roof["achieved"] / roof["peak"]

    # ---- end to end through the public API with host buffers
    host = {k: torch.from_numpy(win[k]).pin_memory() for k in
            ("poses", "disps", "disps_sens", "targets", "weights", "eta", "ii", "jj", "intrinsics")}
    h2d = sum(host[k].numel() * host[k].element_size() for k in host)

    hp0, hd0 = host["poses"].clone(), host["disps"].clone()
    d2h = (t1 - t0) * 7 * 4 + host["disps"].numel() * 4 + (t1 - t0) * 6 * 4

    def e2e_step():
        # the host-buffer entry point: inputs copied from pinned memory, iterations on the device, updated poses,
        # disparities and dx copied back, stream synchronised before it returns
        host["poses"].copy_(hp0)
        host["disps"].copy_(hd0)
        db.ba_host(host["poses"], host["disps"], host["intrinsics"], host["disps_sens"], host["targets"],
                   host["weights"], host["eta"], host["ii"], host["jj"], t0, t1, iters, 1e-4, 0.1, False)

    for _ in range(args.warmup):
        e2e_step()
    e2e_ms = []
    for _ in range(args.steps):
        flush()
        torch.cuda.synchronize()
        a = time.perf_counter()
        e2e_step()
        e2e_ms.append((time.perf_counter() - a) * 1e3)
    e2e_val = statistics.mean(e2e_ms) / iters

    # ---- correlation half of the metric: volume + pyramid build, 4-level lookup (same 48 edges)
    corr = bench_corr(db, synth, cfg, win, dev, flush, peaks, args)

    # ---- the reference's CUDA `ba` on the same window, same GPU, same run (wall clock: it synchronises
    #      with the host ~20 times per iteration)
    ref_cuda = None
    try:
        import build_ref
        rmod = build_ref.load()
        if rmod is not None:
            tr = window_tensors(win, dev)
            rp0, rd0 = tr["poses"].clone(), tr["disps"].clone()
            rms = []
            for it in range(3 + min(args.steps, 20)):
                tr["poses"].copy_(rp0)
                tr["disps"].copy_(rd0)
                flush()
                torch.cuda.synchronize()
                a = time.perf_counter()
                rmod.ba(*ba_call_args(tr), t0, t1, iters, 1e-4, 0.1, False)
                torch.cuda.synchronize()
                if it >= 3:
                    rms.append((time.perf_counter() - a) * 1e3)
            ref_cuda = dict(ba_ms_per_iter=statistics.mean(rms) / iters, unit="ms/iter",
                            note="unmodified reference droid_backends.ba (oracle/_ref) on this GPU")
    except Exception as ex:
        ref_cuda = dict(error=str(ex)[:200])

    line = dict(metric="dense-BA ms/iter", value=step_ms / iters, unit="ms/iter", n_gpus=world, steps=args.steps,
                warmup=args.warmup, ms_per_step=step_ms, higher_is_better=False, scaling="weak",
                vs_baseline=None, dtype="f32", data="synthetic",
                config=frontend_config(cfg, win["poses"].shape[0], E, ht, wd, iters), replicas=world,
                roofline=roof, stages_ms=stage,
                e2e=dict(value=e2e_val, unit="ms/iter", h2d_bytes_per_step=h2d, d2h_bytes_per_step=d2h,
                         ms_per_step=statistics.mean(e2e_ms)),
                gpu_launches=args.steps * (1 + iters * 4), corr=corr, reference_cuda_same_gpu=ref_cuda)
    return line, clk


def bench_corr(db, synth, cfg, win, dev, flush, peaks, args):
    ht, wd, P, E, _ = synth.CONFIGS[cfg]
    HW = ht * wd
    feats = torch.from_numpy(synth.make_features(cfg, seed=0, n_frames=win["poses"].shape[0])).to(dev)
    ii, jj = torch.from_numpy(win["ii"]).to(dev), torch.from_numpy(win["jj"]).to(dev)
    f1, f2 = feats[ii].contiguous(), feats[jj].contiguous()
    coords = torch.from_numpy(synth.lookup_coords(win, 0.05, 0)).to(dev)
    n = f1.shape[0]
    steps = max(3, min(args.steps, 10))
    pyr = db.corr_volume_pyramid(f1, f2, 4)
    vol_ms = statistics.mean(timed_steps(lambda: db.corr_volume_pyramid(f1, f2, 4), steps, 3, flush))
    # the call the pipeline makes (dbaf/covisible_graph.py:129-131): per-frame maps + edge lists, no gather
    idx_ms = statistics.mean(timed_steps(lambda: db.corr_volume_pyramid_indexed(feats, feats, ii, jj, 4), steps, 3, flush))
    wpeak, wpeak_kinds = write_only_peak(dev)
    look_ms = statistics.mean(timed_steps(lambda: db.corr_lookup_pyramid(pyr, coords, 3), steps, 3, flush))
    # the reference's own path for the same work, on this GPU, in this run: torch.matmul (cuBLAS) + 3x
    # avg_pool2d (dbaf/modules/corr.py:24-38,63-71) and its corr_index_forward kernel (oracle/_ref)
    C = f1.shape[1]

    def ref_volume():
        a = f1.reshape(n, C, HW) / 4.0
        b = f2.reshape(n, C, HW) / 4.0
        c = torch.matmul(a.transpose(1, 2), b).reshape(n * HW, 1, ht, wd)
        out = []
        for i in range(4):
            out.append(c.view(n, ht, wd, ht // 2 ** i, wd // 2 ** i))
            c = torch.nn.functional.avg_pool2d(c, 2, stride=2)
        return out

    ref = dict(volume_pyramid_ms=statistics.mean(timed_steps(ref_volume, steps, 3, flush)))

    def ref_gather_volume():
        nonlocal f1, f2
        f1, f2 = feats[ii], feats[jj]
        return ref_volume()
    ref["gather_volume_pyramid_ms"] = statistics.mean(timed_steps(ref_gather_volume, steps, 3, flush))
    f1, f2 = feats[ii].contiguous(), feats[jj].contiguous()
    try:
        import build_ref
        rmod = build_ref.load()
        if rmod is not None:
            ref["lookup_ms"] = statistics.mean(timed_steps(
                lambda: [rmod.corr_index_forward(pyr[l], (coords / 2 ** l).contiguous(), 3) for l in range(4)],
                steps, 3, flush))
    except Exception as ex:      # the reference extension is optional on the box
        ref["error"] = str(ex)[:200]
    # the caller that feeds the lookup: DepthVideo.reproject as one kernel (4 B read + 12 B written per pixel-edge)
    tw = window_tensors(win, dev)
    rep_ms = statistics.mean(timed_steps(lambda: db.reproject(tw["poses"], tw["disps"], tw["intrinsics"], ii, jj), steps, 3, flush))
    rep_bytes = n * HW * 16
    lvl = sum((ht >> l) * (wd >> l) for l in range(4))
    vol_bytes = n * (2 * HW * 128 * 2 + 2 * HW * lvl)             # read both fmaps, write 4 levels
    vol_flop = n * 2.0 * HW * HW * 128
    look_bytes = n * HW * 912                                     # 64 taps + 49 outputs per level + coords
    idx_bytes = n * 2 * HW * lvl + 2 * feats.numel() * 2          # write 4 levels; read + write the frame maps once
    return dict(
        volume_pyramid_indexed=dict(
            edges_per_s=n / (idx_ms * 1e-3), ms=idx_ms,
            roofline=dict(bound="hbm", kernel="k_volume_tc", achieved=idx_bytes / (idx_ms * 1e-3) / 1e9, peak=peaks["hbm"],
                          unit="GB/s", frac=idx_bytes / (idx_ms * 1e-3) / 1e9 / peaks["hbm"],
                          traffic=NCU_TRAFFIC_BYTES.get((cfg, "k_volume_tc")),
                          write_only_peak=wpeak, write_only_peak_by=wpeak_kinds,
                          frac_of_write_only_peak=idx_bytes / (idx_ms * 1e-3) / 1e9 / wpeak,
                          note="97 % of the bytes are writes; write_only_peak = best of cudaMemsetAsync and a 128-bit "
                               "store kernel over 2 GiB, measured in this run")),
        volume_pyramid=dict(edges_per_s=n / (vol_ms * 1e-3), ms=vol_ms,
                            roofline=dict(bound="hbm", achieved=vol_bytes / (vol_ms * 1e-3) / 1e9, peak=peaks["hbm"],
                                          unit="GB/s", frac=vol_bytes / (vol_ms * 1e-3) / 1e9 / peaks["hbm"],
                                          write_only_peak=wpeak),
                            tensor=dict(achieved=vol_flop / (vol_ms * 1e-3) / 1e12, peak=peaks["tf_burst"], unit="TFLOP/s",
                                        frac=vol_flop / (vol_ms * 1e-3) / 1e12 / peaks["tf_burst"])),
        lookup=dict(edges_per_s=n / (look_ms * 1e-3), ms=look_ms,
                    roofline=dict(bound="hbm", kernel="k_corr_lookup", achieved=look_bytes / (look_ms * 1e-3) / 1e9,
                                  peak=peaks["hbm"], unit="GB/s", frac=look_bytes / (look_ms * 1e-3) / 1e9 / peaks["hbm"],
                                  traffic=NCU_TRAFFIC_BYTES.get((cfg, "k_corr_lookup")))),
        reproject=dict(ms=rep_ms, edges_per_s=n / (rep_ms * 1e-3),
                       roofline=dict(bound="hbm", kernel="k_reproject", achieved=rep_bytes / (rep_ms * 1e-3) / 1e9,
                                     peak=peaks["hbm"], unit="GB/s", frac=rep_bytes / (rep_ms * 1e-3) / 1e9 / peaks["hbm"]),
                       note="reference: ~10 torch elementwise ops + lietorch act4 (not installed here), not timed"),
        edges=n, reference_same_gpu=ref)


def cpu_baseline(cfg, budget_s=20.0):
    """Oracle port (numpy, float32, one core) of `ba` on the host cores of this box.  Frontend windows are run whole;
    the global-BA workload is sampled: the first 48 keyframes of the same banded graph (same +-10 neighbourhood, same
    frame size), scaled by the edge count -- everything upstream of the solve is linear in the edges, and the
    sample's dense solve (288 unknowns) is cheaper per pose than the full one, so the scaled figure flatters the CPU."""
    import dbaf_oracle as oracle
    from dbaf_b200 import synth
    try:
        from threadpoolctl import threadpool_limits
    except Exception:
        threadpool_limits = None
    big = cfg == "c5"
    if big:
        full_edges = len(synth.banded_graph(synth.CONFIGS["c5"][2], reach=10)[0])
        win = synth.make_window("g24", seed=0, graph="banded", reach=10, n_extra_frames=24, pose_noise=0.005)
        iters, max_runs = 1, 3
    else:
        win = synth.make_window(cfg, seed=0)
        iters, max_runs = 2, 6
    times = []
    t_start = time.perf_counter()

    def one():
        p, d = win["poses"].copy(), win["disps"].copy()
        a = time.perf_counter()
        oracle.ba(p, d, win["intrinsics"], win["disps_sens"], win["targets"], win["weights"], win["eta"],
                  win["ii"], win["jj"], win["t0"], win["t1"], iters, 1e-4, 0.1, False, dtype=np.float32)
        times.append((time.perf_counter() - a) * 1e3)

    ctx = threadpool_limits(limits=1) if threadpool_limits else None
    try:
        if ctx:
            ctx.__enter__()
        one()
        while time.perf_counter() - t_start < budget_s and len(times) < max_runs:
            one()
    finally:
        if ctx:
            ctx.__exit__(None, None, None)
    best = statistics.median(times) / iters
    if big:
        scale = full_edges / len(win["ii"])
        return dict(value=best * scale, unit="ms/iter", cores=1, kind="port", measured_ms=best, scale=scale,
                    sample=f"{len(times)} x oracle.ba (numpy float32, 1 iteration) on a {win['poses'].shape[0]}-keyframe, "
                           f"{len(win['ii'])}-edge window of the same banded graph at 64x64, scaled by "
                           f"{full_edges}/{len(win['ii'])} edges")
    return dict(value=best, unit="ms/iter", cores=1, kind="port",
                sample=f"{len(times)} x oracle.ba (numpy float32, 2 iterations) on the full {cfg} window")


# --------------------------------------------------------------------- IMU-mode call sequence
def bench_bacore(db, synth, cfg, dev, flush, steps):
    """What DBA-Fusion's IMU mode actually calls (dbaf/depth_video.py:392-399, 524-558): BACore.init, then twice
    {hessian -> float64 H, v on the host -> (GTSAM) -> retract(dx from the host)}, plus the marginalisation call
    (a second BACore with zeroed sensor depths, one hessian, H[i,i] += 0.00025).  GTSAM itself is outside the path;
    dx is a fixed small vector.  Wall clock (every call ends in a host synchronisation), ours and the reference's."""
    win = synth.make_window(cfg, seed=0, sensor_frac=0.3)
    t0, t1 = win["t0"], win["t1"]
    n = 6 * (t1 - t0)
    dx_host = torch.from_numpy(np.random.default_rng(3).normal(0, 1e-3, n))
    out = {}
    mods = [("ours", db)]
    try:
        import build_ref
        rmod = build_ref.load()
        if rmod is not None:
            mods.append(("reference", rmod))
    except Exception:
        pass
    for name, mod in mods:
        t = window_tensors(win, dev)
        p0, d0 = t["poses"].clone(), t["disps"].clone()
        zs = torch.zeros_like(t["disps_sens"])
        H = torch.zeros(n, n, dtype=torch.float64)
        v = torch.zeros(n, dtype=torch.float64)

        def update():
            core = mod.BACore()
            core.init(*ba_call_args(t), t0, t1, 2, 1e-4, 0.1, False)
            for _ in range(2):
                core.hessian(H, v)
                core.retract(dx_host)
            del core

        def marginalise():
            core = mod.BACore()
            core.init(t["poses"], t["disps"], t["intrinsics"], zs, t["targets"], t["weights"], t["eta"], t["ii"],
                      t["jj"], t0, t1, 2, 1e-4, 0.1, False)
            core.hessian(H, v)
            for i in range(6):
                H[i, i] += 0.00025
            del core

        rec = {}
        for key, fn in (("update_ms", update), ("marginalise_ms", marginalise)):
            ts = []
            for it in range(3 + steps):
                t["poses"].copy_(p0)
                t["disps"].copy_(d0)
                flush()
                torch.cuda.synchronize()
                a = time.perf_counter()
                fn()
                torch.cuda.synchronize()
                if it >= 3:
                    ts.append((time.perf_counter() - a) * 1e3)
            rec[key] = statistics.mean(ts)
        out[name] = rec
    out["d2h_bytes_per_hessian"] = (n * n + n) * 8
    out["h2d_bytes_per_retract"] = n * 4
    out["workload"] = f"{cfg}: BACore.init + 2 x (hessian -> host, retract) / marginalisation hessian, {n}x{n} float64 H"
    return out


def bench_proximity(db, synth, dev, steps):
    """edge selection for a new keyframe (dbaf_frontend.py:241: t0 = t-5, t1 = t-window) from the
    DepthVideo buffers to the filtered edge list -- two frame_distance launches, the selection kernel, the duplicate
    filter, one host read of the count -- against the numpy restatement of the reference's Python loops fed with the
    same distances (the reference itself synchronises once per candidate and needs lietorch to import)."""
    import dbaf_oracle as oracle
    from dbaf_b200 import graph
    win = synth.make_window("c2", seed=3)
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    poses, disps, intr = T(win["poses"]), T(win["disps"]), T(win["intrinsics"])
    t = win["poses"].shape[0]
    t0, t1 = t - 5, max(t - 12, 0)
    act, bad, inac = (T(win["ii"][:20]), T(win["jj"][:20])), (T(win["ii"][20:24]), T(win["jj"][20:24])), (T(win["ii"][24:30]), T(win["jj"][24:30]))
    kw = dict(rad=2, nms=1, beta=0.3, thresh=60.0, stereo=False, max_factors=48, skip_edge=[-4, -5, -6], frontend_window=5)

    def ours():
        return graph.proximity_factors(poses, disps, intr, t, t0, t1, act, bad, inac, **kw)

    def port():
        ci, cj, cc = oracle.proximity_candidates(t0, t1, t, kw["skip_edge"], kw["frontend_window"])
        d = (0.5 * (db.frame_distance(poses, disps, intr, T(ci), T(cj), 0.3)
                    + db.frame_distance(poses, disps, intr, T(cj), T(ci), 0.3))).cpu().numpy()
        ei, ej = oracle.proximity_edges(d, ci, cj, cc, win["ii"][:30], win["jj"][:30], t0, t1, t, 2, 1, 60.0, False, 48)
        return oracle.filter_repeated_edges(ei, ej, np.concatenate([win["ii"][:20], win["ii"][24:30]]),
                                            np.concatenate([win["jj"][:20], win["jj"][24:30]]))

    out = {}
    for name, fn in (("ours_ms", ours), ("numpy_port_ms", port)):
        ts = []
        for it in range(3 + steps):
            torch.cuda.synchronize()
            a = time.perf_counter()
            r = fn()
            torch.cuda.synchronize()
            if it >= 3:
                ts.append((time.perf_counter() - a) * 1e3)
        out[name] = statistics.mean(ts)
        out["edges_selected"] = int(len(r[0]))
    out["workload"] = f"c2 buffers, t = {t}, t0 = {t0}, t1 = {t1}: {(t - t0) * (t - t1)} grid candidates + 3 skip edges, 30 existing edges"
    out["host_syncs"] = dict(ours=1, reference="one .item() per candidate, per stamp and per edge (dbaf/covisible_graph.py:61-72,384-431)")
    return out


def bench_altcorr(db, synth, dev, flush, steps):
    """BASELINE.json configs[3]: the on-the-fly correlation (altcorr_forward over the 4 feature-pooled levels,
    dbaf/modules/corr.py:91-139) against the materialised volume + lookup on the same 36-edge window."""
    cfg = "c4"
    ht, wd, P, E, _ = synth.CONFIGS[cfg]
    win = synth.make_window(cfg, seed=0)
    feats = torch.from_numpy(synth.make_features(cfg, seed=0, n_frames=win["poses"].shape[0])).to(dev)
    ii, jj = torch.from_numpy(win["ii"]).to(dev), torch.from_numpy(win["jj"]).to(dev)
    n = len(win["ii"])
    coords0 = torch.from_numpy(synth.lookup_coords(win, 0.05, 0)).to(dev)          # [E,2,ht,wd]
    coords = coords0.permute(0, 2, 3, 1)[:, None].contiguous()                     # [E,1,ht,wd,2]
    f1 = (feats[ii].float() / 4.0).permute(0, 2, 3, 1).contiguous()
    f2 = feats[jj].float() / 4.0
    pyr2 = [f2.permute(0, 2, 3, 1).contiguous()]
    for _ in range(3):
        f2 = torch.nn.functional.avg_pool2d(f2, 2, stride=2)
        pyr2.append(f2.permute(0, 2, 3, 1).contiguous())

    def alt(mod):
        return [mod.altcorr_forward(f1, pyr2[l], (coords / 2 ** l).contiguous(), 3)[0] for l in range(4)]

    ours_ms = statistics.mean(timed_steps(lambda: alt(db), steps, 3, flush))
    HW = ht * wd
    flop = n * 2.0 * 128 * 64 * 4 * HW
    bytes_ = n * (HW * 128 * 4 + sum((ht >> l) * (wd >> l) for l in range(4)) * 128 * 4 + HW * 196 * 4 + HW * 8)
    out = dict(edges=n, ours_ms=ours_ms, ours_edges_per_s=n / (ours_ms * 1e-3),
               roofline=dict(bound="hbm", kernel="k_altcorr_fwd", achieved=bytes_ / (ours_ms * 1e-3) / 1e9,
                             unit="GB/s", algorithmic_bytes=bytes_, fp32_tflops=flop / (ours_ms * 1e-3) / 1e12))
    pyr = db.corr_volume_pyramid_indexed(feats, feats, ii, jj, 4)
    vol_ms = statistics.mean(timed_steps(lambda: db.corr_volume_pyramid_indexed(feats, feats, ii, jj, 4), steps, 3, flush))
    look_ms = statistics.mean(timed_steps(lambda: db.corr_lookup_pyramid(pyr, coords0, 3), steps, 3, flush))
    out["volume_plus_lookup_ms"] = dict(volume_pyramid=vol_ms, lookup=look_ms)
    try:
        import build_ref
        rmod = build_ref.load()
        if rmod is not None:
            out["reference_ms"] = statistics.mean(timed_steps(lambda: alt(rmod), steps, 3, flush))
    except Exception as ex:
        out["reference_error"] = str(ex)[:200]
    return out


# --------------------------------------------------------------------- reference arm
def bench_reference(args, dev):
    """The unmodified reference on the same workload (oracle/_ref when it is built, else the oracle port).
    The sample is bounded to about two minutes of wall clock: the reference's host-side Schur loop makes
    one global-BA iteration take tens of seconds, so there the sample is a single iteration."""
    import build_ref
    from dbaf_b200 import synth
    cfg = args.workload
    ht, wd, P, E, _ = synth.CONFIGS[cfg]
    big = cfg == "c5"
    win = synth.make_window(cfg, seed=0, graph="banded", pose_noise=0.005) if big else synth.make_window(cfg, seed=0)
    n_edges = len(win["ii"])
    ref = build_ref.load() if torch.cuda.is_available() else None
    iters = 1 if big else 2
    if ref is None:
        cb = cpu_baseline(cfg, budget_s=60.0)
        val, step_ms, kind, n_steps = cb["value"], cb["value"] * iters, cb, 0
    else:
        t = window_tensors(win, dev)
        p0, d0 = t["poses"].clone(), t["disps"].clone()
        flush = L2Flusher(dev)

        def one():
            t["poses"].copy_(p0)
            t["disps"].copy_(d0)
            flush()
            torch.cuda.synchronize()
            a = time.perf_counter()
            ref.ba(*ba_call_args(t), win["t0"], win["t1"], iters, 1e-4, 0.1, False)
            torch.cuda.synchronize()
            return (time.perf_counter() - a) * 1e3          # wall clock: the reference syncs with the host

        budget_ms = 120e3
        first = one()
        ms = [first]
        if first < 5e3:                                       # cheap enough: proper warm-up, then K steps
            for _ in range(args.warmup - 1):
                one()
            ms = []
            while len(ms) < args.steps and sum(ms) + first < budget_ms:
                ms.append(one())
        n_steps = len(ms)
        step_ms = statistics.mean(ms)
        val = step_ms / iters
        kind = dict(value=val, unit="ms/iter", cores=1, kind="reference",
                    sample=f"{n_steps} x reference droid_backends.ba (oracle/_ref: CUDA kernels + host Eigen float64 "
                           f"assembly/solve on one core), {iters} iteration(s) each, full {cfg} window"
                           + (" (single cold call: one call takes tens of seconds)" if first >= 5e3 else ""))
    return dict(impl="reference", metric="dense-BA ms/iter", value=val, unit="ms/iter", n_gpus=1, steps=n_steps,
                warmup=args.warmup, ms_per_step=step_ms, higher_is_better=False,
                scaling="strong" if big else "weak", vs_baseline=None, dtype="f32", data="synthetic",
                config=(global_ba_config() if big else frontend_config(cfg, win["poses"].shape[0], n_edges, ht, wd, iters)),
                cpu_baseline=kind,
                e2e=dict(value=val, unit="ms/iter", h2d_bytes_per_step=0, d2h_bytes_per_step=0))


# ----------------------------------------------------------------------------- main
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sub", action="store_true", help="skip the frontend / BACore / altcorr sub-records at N = 1")
    args = ap.parse_args()
    args.warmup = max(3, args.warmup)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if rank != 0:
            return
        args.workload = args.workload or "c5"
        dev = torch.device("cuda:0") if torch.cuda.is_available() else torch.device("cpu")
        print(json.dumps(bench_reference(args, dev)))
        return

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback); use --impl reference for the CPU port")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.distributed.init_process_group("nccl", device_id=dev)

    default_run = args.workload is None
    args.workload = args.workload or "c5"
    if args.workload.startswith("c5"):
        from dbaf_b200 import global_ba
        line, clk = global_ba.bench(args, dev, rank, world, ClockSampler, L2Flusher, measured_peaks())
        if default_run and world == 1 and not args.no_sub:
            # the frontend halves of the metric, as named sub-records of the same line
            import copy
            import droid_backends as db
            from dbaf_b200 import synth
            fa = copy.copy(args)
            fa.steps = min(args.steps, 20)
            for cfg in ("c2", "c2_real", "c3"):
                fa.workload = cfg
                fl, _ = bench_local(fa, dev, rank, world)
                line[f"frontend_{cfg}"] = {k: fl[k] for k in ("value", "unit", "ms_per_step", "config", "stages_ms",
                                                               "roofline", "e2e", "gpu_launches", "corr",
                                                               "reference_cuda_same_gpu")}
            flush = L2Flusher(dev)
            line["bacore"] = {cfg: bench_bacore(db, synth, cfg, dev, flush, min(args.steps, 10)) for cfg in ("c2_real", "c3")}
            line["c4_altcorr"] = bench_altcorr(db, synth, dev, flush, max(3, min(args.steps, 10)))
            line["proximity"] = bench_proximity(db, synth, dev, max(3, min(args.steps, 10)))
    else:
        line, clk = bench_local(args, dev, rank, world)

    if world > 1:
        v = torch.tensor([line["ms_per_step"]], device=dev, dtype=torch.float64)
        torch.distributed.all_reduce(v, op=torch.distributed.ReduceOp.MAX)
        scale = float(v.item()) / line["ms_per_step"]
        line["ms_per_step"] = float(v.item())
        line["value"] *= scale
    rc = 0
    if rank == 0:
        line["clocks"] = clk.summary()
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline(args.workload)
        print(json.dumps(line))
        par = line.get("parity_vs_1gpu")
        if par and max(par["dx_rel"], par["poses_rel"], par["disps_rel"]) > par["bound"]:
            print(f"bench.py: the sharded iteration differs from the single-rank one: {par}", file=sys.stderr)
            rc = 3
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()
    if rc:
        sys.exit(rc)


if __name__ == "__main__":
    main()
