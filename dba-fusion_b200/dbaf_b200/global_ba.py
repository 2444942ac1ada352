"""Edge-sharded global bundle adjustment over `torch.distributed` (one process per GPU).

Everything upstream of the pose solve is a sum over edges grouped by their SOURCE frame
(SURVEY.md 8e): the Jacobian kernel reads only the source frame's disparity, and C_k, w_k and every
Schur term of depth frame k come from the edges leaving k.  So each rank owns a contiguous range of
depth frames together with all their out-edges, linearises them into its own copy of the reduced
camera system, and ONE all-reduce (sum, float64, in place on the contiguous H|v span of the
workspace) gives every rank the identical full system.  The solve is replicated (bit-identical dx on
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

    def linearize(self, alpha, k_lo, k_hi):
        self.w.linearize(alpha, zero_first=True, k_lo=k_lo, k_hi=k_hi)

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
        e.linearize(alpha, self.k_lo, self.k_hi)
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
        idx = self.band_index(sys_flat.device)
        if idx is None:
            dist.all_reduce(sys_flat, op=dist.ReduceOp.SUM, group=self.group)
            return
        packed = sys_flat.index_select(0, idx)
        dist.all_reduce(packed, op=dist.ReduceOp.SUM, group=self.group)
        sys_flat.index_copy_(0, idx, packed)

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


# --------------------------------------------------------------------------- benchmark arm
def bench(args, dev, rank, world, ClockSampler, L2Flusher, peaks):
    """bench.py arm for N > 1 (and --workload c5): configs[4], 1000 keyframes, +-10 neighbours."""
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
    for name, fn in (("linearize+schur", lambda: eng.linearize(0.05, gba.k_lo, gba.k_hi)),
                     ("allreduce", (lambda: gba.reduce_system(eng.system())) if world > 1 else (lambda: None)),
                     ("solve", lambda: eng.solve(1e-4, 0.1, gba.band)),
                     ("retract", lambda: eng.retract(gba.k_lo, gba.k_hi))):
        ts = []
        for _ in range(3):
            reset()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            b.synchronize()
            ts.append(a.elapsed_time(b))
        stages[name] = min(ts)
    HW = ht * wd
    K = len(gba.kx)
    lin_bytes = n_edges * HW * 20 + K * HW * 12
    roof = dict(bound="hbm", kernel="k_linearize2 + k_schur_tc (rank 0 shard)", unit="GB/s", peak=peaks["hbm"],
                achieved=lin_bytes / world / (stages["linearize+schur"] * 1e-3) / 1e9, traffic=None,
                peak_source=peaks["src"], algorithmic_bytes=lin_bytes // world, ms=stages["linearize+schur"])
    roof["frac"] = roof["achieved"] / roof["peak"]
    line = dict(metric="dense-BA ms/iter", value=step_ms, unit="ms/iter", n_gpus=world, steps=args.steps,
                warmup=max(3, args.warmup), ms_per_step=step_ms, higher_is_better=False, scaling="strong",
                vs_baseline=None, dtype="f32", data="synthetic",
                config=dict(workload=f"c5 global BA: {P} keyframes, {n_edges} edges (+-10 neighbours), {ht}x{wd} "
                                     f"disparity maps, 1 GN iteration per step, depth frames sharded over {world} GPU(s)",
                            l2="512 MB read+write between steps", band_blocks=gba.band,
                            shard=[list(r) for r in gba.ranges]),
                roofline=roof, stages_ms=stages,
                e2e=dict(value=step_ms, unit="ms/iter", h2d_bytes_per_step=0, d2h_bytes_per_step=0,
                         note="global BA state lives on the GPUs between iterations; no host traffic per step"),
                gpu_launches=args.steps * 9)   # linearize2, schur_tc, schur_smem, schur, 3 x banded solve, memset-free retract
    return line, clk
