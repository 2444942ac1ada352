"""Edge-sharded global bundle adjustment over `torch.distributed` (one process per GPU).

Everything upstream of the pose solve is a sum over edges grouped by their SOURCE frame
(SURVEY.md 8e): the Jacobian kernel reads only the source frame's disparity, and C_k, w_k and every
Schur term of depth frame k come from the edges leaving k.  So each rank owns a contiguous range of
depth frames together with all their out-edges, linearises them into its own copy of the reduced
camera system, and ONE all-reduce (sum, float64; of the band of H and v packed into one dense strip
when the frame graph is banded, else in place on the contiguous H|v span of the workspace) gives every
rank the identical full system.  The solve is replicated (bit-identical dx on
every rank, no broadcast), poses are updated everywhere, disparities only by their owner -- no
per-pixel data ever crosses GPUs.

The reference has no multi-GPU path; on one GPU it assembles and solves this system on the host
(src/droid_kernels.cu:1162-1391).
"""
from __future__ import annotations

import json
import statistics
import time

import numpy as np
import torch
import torch.distributed as dist


# --------------------------------------------------------------------------- host logic
def depth_frames(ii, t0, t1):
    """kx = sorted unique of cat(arange(t0,t1), ii)  (src/droid_kernels.cu:1416-1424), on the host."""
    return np.unique(np.concatenate([np.arange(t0, t1, dtype=np.int64), np.asarray(ii, np.int64)]))


def shard_ranges(ii, t0, t1, world):
    """Contiguous ranges [k_lo, k_hi) of depth-frame positions, one per rank, balanced by the work a
    frame causes: its out-edges (Jacobians) plus the square of its row count (Schur pairs)."""
    kx = depth_frames(ii, t0, t1)
    pos = {int(f): n for n, f in enumerate(kx)}
    deg = np.zeros(len(kx), np.int64)
    for f in np.asarray(ii, np.int64):
        deg[pos[int(f)]] += 1
    cost = 1.0 + deg + (1.0 + deg) ** 2 / 8.0
    cum = np.concatenate([[0.0], np.cumsum(cost)])
    bounds = [0]
    for r in range(1, world):
        target = cum[-1] * r / world
        bounds.append(int(np.searchsorted(cum, target, side="left")))
    bounds.append(len(kx))
    bounds = np.maximum.accumulate(np.array(bounds))
    return [(int(bounds[r]), int(bounds[r + 1])) for r in range(world)]


def band_blocks(ii, jj, t0, t1):
    """Block half-bandwidth of the reduced camera system: poses a, b are coupled when they share a depth
    frame k, i.e. a, b in {k} + targets(k); the largest |a - b| over all frames is the band."""
    lo, hi = {}, {}

    def add(k, f):
        if t0 <= f < t1:
            lo[k] = min(lo.get(k, f), f)
            hi[k] = max(hi.get(k, f), f)

    for i, j in zip(np.asarray(ii, np.int64).tolist(), np.asarray(jj, np.int64).tolist()):
        add(i, i)
        add(i, j)
    return max([hi[k] - lo[k] for k in lo] + [0])


class CudaEngine:
    """The staged C ABI (dbaf_ba_prepare / linearize / solve_banded / retract) on this rank's GPU."""

    def __init__(self, poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1):
        import droid_backends as db
        self.db = db
        self.w = db._Window(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1)
        self.w.prepare()

    def linearize(self, alpha, k_lo, k_hi, band=0):
        # every block of the reduced system lies within `band` pose blocks of the diagonal: clear only that
        self.w.zero_system(band)
        self.w.linearize(alpha, zero_first=False, k_lo=k_lo, k_hi=k_hi)

    def system(self):
        return self.w.Hv()

    def solve(self, lm, ep, band):
        self.w.solve(lm, ep, band=band)

    def retract(self, k_lo, k_hi):
        self.w.retract(update_poses=True, update_disps=True, k_lo=k_lo, k_hi=k_hi, want_dx=False, want_dz=False)


class GlobalBA:
    """ba(...) for a large frame graph, sharded over the ranks of `group`."""

    def __init__(self, engine, ii_host, jj_host, t0, t1, group=None):
        self.engine = engine
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.ranges = shard_ranges(ii_host, t0, t1, self.world)
        self.k_lo, self.k_hi = self.ranges[self.rank]
        self.band = band_blocks(ii_host, jj_host, t0, t1)
        self.kx = depth_frames(ii_host, t0, t1)
        self.P = int(t1) - int(t0)
        self._band_idx = None

    def iterate(self, lm=1e-4, ep=0.1, alpha=0.05):
        e = self.engine
        e.linearize(alpha, self.k_lo, self.k_hi, self.band)
        if self.world > 1:
            self.reduce_system(e.system())
        e.solve(lm, ep, self.band)
        e.retract(self.k_lo, self.k_hi)

    def band_index(self, device):
        """Flat indices of the entries of H|v that can be non-zero: pose blocks at most `band` apart, plus v.
        None when the band covers most of the matrix anyway."""
        P, B = self.P, self.band
        if B <= 0 or 2 * B + 1 >= 0.6 * P:
            return None
        if self._band_idx is None or self._band_idx.device != device:
            blk = torch.arange(P, device=device)
            keep = ((blk[:, None] - blk[None, :]).abs() <= B).repeat_interleave(6, 0).repeat_interleave(6, 1)
            n = 6 * P
            self._band_idx = torch.cat([keep.reshape(-1).nonzero().squeeze(1),
                                        n * n + torch.arange(n, device=device)])
        return self._band_idx

    def reduce_system(self, sys_flat):
        """Sum the reduced camera system over the ranks (the one exchange step of the path).  Every rank's
        contribution lies inside the band of the frame graph, so only the band travels: 12 MB instead of the
        288 MB of the dense 6000 x 6000 float64 H at the 1000-keyframe size."""
        w = getattr(self.engine, "w", None)
        if w is not None and hasattr(w, "band_strip") and 0 < self.band and 2 * self.band + 1 < 0.6 * self.P:
            # the CUDA engine packs / unpacks the band with its own kernels (dbaf_ba_band_strip)
            self._strip = w.band_strip(self.band, getattr(self, "_strip", None), pack=True)
            dist.all_reduce(self._strip, op=dist.ReduceOp.SUM, group=self.group)
            w.band_strip(self.band, self._strip, pack=False)
            return
        idx = self.band_index(sys_flat.device)
        if idx is None:
            dist.all_reduce(sys_flat, op=dist.ReduceOp.SUM, group=self.group)
            return
        packed = sys_flat.index_select(0, idx)
        dist.all_reduce(packed, op=dist.ReduceOp.SUM, group=self.group)
        sys_flat.index_copy_(0, idx, packed)

    def launches_per_iteration(self):
        """Kernels of this library enqueued by one `iterate` on a dense graph: k_zero_band, k_depth_diag, k_linschur_tc, then
        k_linearize2 / k_schur_smem / k_schur for the frames those left, the banded solve's kernels (sweeps, spike,
        separator chain, back-substitutions), k_backsub_fused, k_backsub_rc."""
        import droid_backends as db
        return 6 + db.banded_solve_launches(self.P, self.band) + 2

    def run(self, iterations, lm=1e-4, ep=0.1):
        for _ in range(iterations):
            self.iterate(lm, ep)

    def gather_disps(self, disps):
        """Make every rank's `disps` complete (owners broadcast their frames); only needed to inspect the
        result -- the iteration itself never moves disparities between GPUs."""
        if self.world == 1:
            return
        owner = torch.zeros(disps.shape[0], dtype=torch.bool, device=disps.device)
        owner[torch.as_tensor(self.kx[self.k_lo:self.k_hi], device=disps.device)] = True
        contrib = torch.where(owner[:, None, None], disps, torch.zeros_like(disps))
        dist.all_reduce(contrib, group=self.group)
        touched = owner.float()
        dist.all_reduce(touched, group=self.group)
        disps.copy_(torch.where(touched[:, None, None] > 0, contrib, disps))


# --------------------------------------------------------------------------- host-resident windows
def _runs(ids):
    """Sorted integer ids -> list of (first, count) runs of consecutive values."""
    ids = np.asarray(ids, np.int64)
    if ids.size == 0:
        return []
    cut = np.nonzero(np.diff(ids) != 1)[0] + 1
    return [(int(r[0]), int(len(r))) for r in np.split(ids, cut)]


class HostWindow:
    """A window that lives in pinned HOST memory (a backend thread that assembles targets / weights on the CPU or
    receives them over IPC), iterated on this rank's GPU: `GlobalBA.iterate_host` uploads what this rank needs
    (all poses and the edge lists, its own depth frames' disparities / eta / sensor depths, its own edges' targets
    and weights), runs one Gauss-Newton iteration, and reads back the pose update, the window poses and its own
    disparities.  The uploads of targets / weights run on a copy stream in `chunks` pieces of depth frames and the
    linearisation of a piece starts as soon as it has landed."""

    KEYS = ("poses", "disps", "disps_sens", "targets", "weights", "eta", "ii", "jj", "intrinsics")

    def __init__(self, win):
        self.h = {k: torch.from_numpy(np.ascontiguousarray(win[k])).pin_memory() for k in self.KEYS}
        self.t0, self.t1 = int(win["t0"]), int(win["t1"])
        P = self.t1 - self.t0
        self.dx = torch.zeros(max(P, 1), 6, dtype=torch.float32).pin_memory()


def plan_host_chunks(ii, kx, k_lo, k_hi, chunks):
    """Split the depth-frame positions [k_lo,k_hi) into `chunks` pieces; for each piece the runs of consecutive
    frame ids (disps / sens rows), of eta rows, and of edge ids whose source frame lies in the piece."""
    ii = np.asarray(ii, np.int64)
    bounds = np.linspace(k_lo, k_hi, max(1, chunks) + 1).round().astype(int)
    out = []
    for a, b in zip(bounds[:-1], bounds[1:]):
        if b <= a:
            continue
        frames = kx[a:b]
        edges = np.nonzero(np.isin(ii, frames))[0]
        out.append(dict(k=(int(a), int(b)), frame_runs=_runs(frames), eta_run=(int(a), int(b - a)),
                        edge_runs=_runs(edges), n_edges=int(len(edges))))
    return out


def iterate_host(gba, host, dev, plan, copy_stream, lm=1e-4, ep=0.1, alpha=0.05):
    """One global-BA iteration from / to pinned host memory on this rank (see HostWindow).  `dev` is the dict of
    device tensors the engine was built on.  Returns (h2d_bytes, d2h_bytes) of this rank."""
    eng, h = gba.engine, host.h
    main = torch.cuda.current_stream()
    h2d = d2h = 0
    ev_small = torch.cuda.Event()
    copy_stream.wait_stream(main)
    with torch.cuda.stream(copy_stream):
        for k in ("poses", "intrinsics", "ii", "jj"):
            dev[k].copy_(h[k], non_blocking=True)
            h2d += h[k].numel() * h[k].element_size()
        ev_small.record(copy_stream)
        events = []
        for ch in plan:
            for f0, cnt in ch["frame_runs"]:
                for k in ("disps", "disps_sens"):
                    dev[k][f0:f0 + cnt].copy_(h[k][f0:f0 + cnt], non_blocking=True)
                    h2d += cnt * h[k][0].numel() * 4
            a, cnt = ch["eta_run"]
            if h["eta"].shape[0] > 1:
                dev["eta"][a:a + cnt].copy_(h["eta"][a:a + cnt], non_blocking=True)
                h2d += cnt * h["eta"][0].numel() * 4
            for e0, cnt in ch["edge_runs"]:
                for k in ("targets", "weights"):
                    dev[k][e0:e0 + cnt].copy_(h[k][e0:e0 + cnt], non_blocking=True)
                    h2d += cnt * h[k][0].numel() * 4
            ev = torch.cuda.Event()
            ev.record(copy_stream)
            events.append(ev)
    main.wait_event(ev_small)
    eng.w.prepare()                                    # the edge lists were just uploaded: rebuild the tables
    for n, (ch, ev) in enumerate(zip(plan, events)):
        main.wait_event(ev)
        eng.w.linearize(alpha, zero_first=(n == 0), k_lo=ch["k"][0], k_hi=ch["k"][1])
    if not plan:
        eng.w.Hv().zero_()
    if gba.world > 1:
        gba.reduce_system(eng.system())
    eng.solve(lm, ep, gba.band)
    eng.w.retract(update_poses=True, update_disps=True, k_lo=gba.k_lo, k_hi=gba.k_hi, want_dx=False, want_dz=False)
    P = host.t1 - host.t0
    h["poses"][host.t0:host.t1].copy_(dev["poses"][host.t0:host.t1], non_blocking=True)
    host.dx[:P].copy_(eng.w.dx(), non_blocking=True)
    d2h += P * 7 * 4 + P * 6 * 4
    for ch in plan:
        for f0, cnt in ch["frame_runs"]:
            h["disps"][f0:f0 + cnt].copy_(dev["disps"][f0:f0 + cnt], non_blocking=True)
            d2h += cnt * h["disps"][0].numel() * 4
    main.synchronize()
    return h2d, d2h


# --------------------------------------------------------------------------- benchmark arm
C5_WORKLOAD = ("c5 global BA: 1000-frame banded keyframe graph (999 window poses, 19890 edges, +-10 neighbours), "
               "64x64 disparity maps, 1 GN iteration per step")
C5_L2 = "512 MB read+write between steps (the 1.4 GB of inputs are also larger than the 126 MB L2)"


# dram__bytes_read.sum + dram__bytes_write.sum of k_depth_diag (0.721 GB) + k_linschur_tc (1.399 GB) for the whole
# window, from the committed `ncu --set full` capture (profiles/r2_ncu_c5.md)
C5_NCU_TRAFFIC_BYTES = 2.120e9


def c5_config():
    """`config` of the bench line; the reference arm prints the identical dict."""
    return dict(workload=C5_WORKLOAD, l2=C5_L2)


def bench(args, dev, rank, world, ClockSampler, L2Flusher, peaks):
    """The bench line of record at every N: BASELINE.json configs[4] (1000 keyframes, +-10 neighbours), the depth
    frames sharded over the `world` ranks.  `value` = device time of one Gauss-Newton iteration with all inputs
    resident in HBM; `e2e` = the same iteration from / to pinned host memory (HostWindow)."""
    from . import synth
    cfg = "c5"
    ht, wd, P, E, _ = synth.CONFIGS[cfg]
    win = synth.make_window(cfg, seed=0, graph="banded", pose_noise=0.005)
    t = {k: (torch.from_numpy(v.copy()).to(dev) if isinstance(v, np.ndarray) else v) for k, v in win.items()}
    n_edges = len(win["ii"])
    eng = CudaEngine(t["poses"], t["disps"], t["intrinsics"], t["disps_sens"], t["targets"], t["weights"],
                     t["eta"], t["ii"], t["jj"], win["t0"], win["t1"])
    gba = GlobalBA(eng, win["ii"], win["jj"], win["t0"], win["t1"])
    p0, d0 = t["poses"].clone(), t["disps"].clone()
    flush = L2Flusher(dev)

    def reset():
        t["poses"].copy_(p0)
        t["disps"].copy_(d0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        reset()
        gba.iterate()
    ms = []
    with ClockSampler(dev.index or 0) as clk:
        barrier()
        for _ in range(args.steps):
            reset()
            flush()
            barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            gba.iterate()
            b.record()
            b.synchronize()
            ms.append(a.elapsed_time(b))
        barrier()
    step_ms = statistics.mean(ms)
    # stage breakdown on this rank
    stages = {}
    for name, fn in (("linearize+schur", lambda: eng.linearize(0.05, gba.k_lo, gba.k_hi, gba.band)),
                     ("allreduce", (lambda: gba.reduce_system(eng.system())) if world > 1 else (lambda: None)),
                     ("solve", lambda: eng.solve(1e-4, 0.1, gba.band)),
                     ("retract", lambda: eng.retract(gba.k_lo, gba.k_hi))):
        ts = []
        for _ in range(3):
            reset()
            flush()
            barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            b.synchronize()
            ts.append(a.elapsed_time(b))
        stages[name] = min(ts)

    # ---- N ranks == 1 rank: one iteration from the initial state, sharded, against the same iteration done by
    #      this rank alone over all depth frames (every rank holds the whole window in this benchmark)
    reset()
    gba.iterate()
    dx_n = eng.w.dx().clone()
    poses_n, disps_n = t["poses"].clone(), t["disps"].clone()
    gba.gather_disps(disps_n)
    reset()
    eng.linearize(0.05, 0, -1)
    eng.solve(1e-4, 0.1, gba.band)
    eng.retract(0, -1)
    rel = lambda a, b: float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp(min=1e-300))
    par = torch.tensor([rel(dx_n, eng.w.dx()), rel(poses_n, t["poses"]), rel(disps_n, t["disps"])],
                       dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(par, op=dist.ReduceOp.MAX)
    parity = dict(dx_rel=float(par[0]), poses_rel=float(par[1]), disps_rel=float(par[2]), bound=1e-6,
                  what="one iteration sharded over the ranks vs the same iteration on one rank, max over ranks")

    # ---- end to end from pinned host memory
    host = HostWindow(win)
    plan = plan_host_chunks(win["ii"], gba.kx, gba.k_lo, gba.k_hi, 6)
    cs = torch.cuda.Stream(device=dev)
    hp0, hd0 = host.h["poses"].clone(), host.h["disps"].clone()
    e2e_ms, h2d, d2h = [], 0, 0
    for it in range(3 + max(3, min(args.steps, 10))):
        host.h["poses"].copy_(hp0)
        host.h["disps"].copy_(hd0)
        flush()
        barrier()
        a = time.perf_counter()
        h2d, d2h = iterate_host(gba, host, t, plan, cs)
        if it >= 3:
            e2e_ms.append((time.perf_counter() - a) * 1e3)
    e2e = torch.tensor([statistics.mean(e2e_ms), float(h2d), float(d2h)], dtype=torch.float64, device=dev)
    if world > 1:
        e2e_max = e2e.clone()
        dist.all_reduce(e2e_max, op=dist.ReduceOp.MAX)
        dist.all_reduce(e2e, op=dist.ReduceOp.SUM)
        e2e[0] = e2e_max[0]
    HW = ht * wd
    K = len(gba.kx)
    lin_bytes = n_edges * HW * 20 + K * HW * 12
    roof = dict(bound="hbm", kernel="k_depth_diag + k_linschur_tc (this rank's shard: linearisation and Schur complement)",
                unit="GB/s", peak=peaks["hbm"],
                achieved=lin_bytes / world / (stages["linearize+schur"] * 1e-3) / 1e9, traffic=C5_NCU_TRAFFIC_BYTES / world,
                peak_source=peaks["src"], algorithmic_bytes=lin_bytes // world, ms=stages["linearize+schur"])
    roof["frac"] = roof["achieved"] / roof["peak"]
    line = dict(metric="dense-BA ms/iter", value=step_ms, unit="ms/iter", n_gpus=world, steps=args.steps,
                warmup=max(3, args.warmup), ms_per_step=step_ms, higher_is_better=False, scaling="strong",
                vs_baseline=None, dtype="f32", data="synthetic", config=c5_config(),
                shard=dict(depth_frame_ranges=[list(r) for r in gba.ranges], band_blocks=gba.band,
                           solve_launches=gba.launches_per_iteration() - 8),
                roofline=roof, stages_ms=stages, parity_vs_1gpu=parity,
                e2e=dict(value=float(e2e[0]), unit="ms/iter", h2d_bytes_per_step=int(e2e[1]),
                         d2h_bytes_per_step=int(e2e[2]), ms_per_step=float(e2e[0]),
                         note="one iteration from / to pinned host memory (global_ba.iterate_host): every rank uploads "
                              "the poses, the edge lists and its own depth frames' inputs, 6 chunks overlapped with "
                              "the linearisation; bytes are summed over the ranks, time is the max"),
                gpu_launches=args.steps * gba.launches_per_iteration())
    return line, clk
