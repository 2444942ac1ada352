"""Edge-sharded global BA: host logic on CPU (gloo, world_size 2, numpy oracle as the per-rank engine)
and the CUDA engine + banded solver on the GPU."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import record, relerr, to_torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class OracleEngine:
    """Per-rank engine for the CPU test: the numpy restatement restricted to a depth-frame range."""

    def __init__(self, oracle, w):
        self.o, self.w = oracle, w
        self.poses, self.disps = w["poses"].astype(np.float64), w["disps"].astype(np.float64)
        self.ii, self.jj, self.t0, self.t1 = w["ii"], w["jj"], w["t0"], w["t1"]
        self.tab = oracle.build_tables(self.ii, self.jj, self.t0, self.t1)
        n = 6 * (self.t1 - self.t0)
        self.Hv = torch.zeros(n * n + n, dtype=torch.float64)

    def linearize(self, alpha, k_lo, k_hi, band=0):
        o, w = self.o, self.w
        frames = self.tab["kx"][k_lo:k_hi]
        mine = np.isin(self.ii, frames)
        Hs, vs, Eii, Eij, Cii, bz = o.jacobian_edges(self.poses, self.disps, w["intrinsics"], w["targets"],
                                                     w["weights"], self.ii, self.jj, np.float64)
        A, b = o.assemble_pose_system(Hs[:, mine], vs[:, mine], self.ii[mine], self.jj[mine], self.t0, self.t1)
        self.C, self.wk, self.Q, self.E = o.depth_system(Cii, bz, Eii, Eij, self.disps, w["disps_sens"], w["eta"],
                                                         self.ii, self.tab, self.t0, self.t1, alpha, np.float64)
        P = self.t1 - self.t0
        S = np.zeros((P, 6, P, 6))
        sv = np.zeros((P, 6))
        for a, bb, n, m, k in o.schur_triples(self.tab, self.t0, self.t1):
            if k_lo <= k < k_hi:
                S[a, :, bb, :] += (self.E[n] * self.Q[k]) @ self.E[m].T
        qw = self.Q * self.wk
        for n in range(self.E.shape[0]):
            k, p = self.tab["kk_exp"][n], self.tab["jj_exp"][n] - self.t0
            if k_lo <= k < k_hi and p >= 0:
                sv[p] += self.E[n] @ qw[k]
        nn = 6 * P
        self.Hv[:nn * nn] = torch.from_numpy((A - S.reshape(nn, nn)).reshape(-1))
        self.Hv[nn * nn:] = torch.from_numpy(b - sv.reshape(-1))

    def system(self):
        return self.Hv

    def solve(self, lm, ep, band):
        nn = 6 * (self.t1 - self.t0)
        H = self.Hv[:nn * nn].numpy().reshape(nn, nn)
        assert band >= 0
        self.dx, _ = self.o.damped_solve(H, self.Hv[nn * nn:].numpy(), lm, ep)

    def retract(self, k_lo, k_hi):
        o = self.o
        dz = o.back_substitute(self.E, self.Q, self.wk, self.dx, self.tab, self.t0, self.t1, np.float64)
        own = np.zeros(len(self.tab["kx"]), bool)
        own[k_lo:k_hi] = True
        o.retract(self.poses, self.disps, self.dx, np.where(own[:, None], dz, 0.0), self.tab, self.t0, self.t1)


def _worker(rank, world, port, seed, out, graph="frontend"):
    for p in (os.path.join(ROOT, "dba-fusion_b200"), os.path.join(ROOT, "oracle")):
        sys.path.insert(0, p)
    import dbaf_oracle as oracle
    from dbaf_b200 import global_ba, synth
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    w = _gloo_window(synth, seed, graph)
    eng = OracleEngine(oracle, w)
    gba = global_ba.GlobalBA(eng, w["ii"], w["jj"], w["t0"], w["t1"])
    # the banded graph exchanges only the band of H; the frontend-style one the whole system
    assert (gba.band_index(torch.device("cpu")) is not None) == (graph == "banded")
    gba.run(2)
    d = torch.from_numpy(eng.disps)
    gba.gather_disps(d)
    if rank == 0:
        np.savez(out, poses=eng.poses, disps=d.numpy(), ranges=np.array(gba.ranges))
    dist.barrier()
    dist.destroy_process_group()


def _gloo_window(synth, seed, graph):
    if graph == "banded":
        return synth.make_window("tiny", seed=seed, n_extra_frames=10, graph="banded", reach=1)
    return synth.make_window("tiny", seed=seed, n_extra_frames=3)


@pytest.mark.parametrize("graph", ["frontend", "banded"])
def test_sharded_global_ba_matches_single_process_gloo(oracle, synth, tmp_path, graph):
    out = str(tmp_path / "res.npz")
    port = 29500 + (os.getpid() % 2000) + (7 if graph == "banded" else 0)
    mp.start_processes(_worker, args=(2, port, 5, out, graph), nprocs=2, join=True, start_method="spawn")
    res = np.load(out)
    w = _gloo_window(synth, 5, graph)
    p, d = w["poses"].astype(np.float64), w["disps"].astype(np.float64)
    oracle.ba(p, d, w["intrinsics"], w["disps_sens"], w["targets"], w["weights"], w["eta"], w["ii"], w["jj"],
              w["t0"], w["t1"], 2, 1e-4, 0.1, False, dtype=np.float64)
    assert res["ranges"].shape == (2, 2) and res["ranges"][0, 1] == res["ranges"][1, 0]
    assert relerr(res["poses"], p) < 1e-9 and relerr(res["disps"], d) < 1e-9


def test_shard_ranges_and_band(oracle, synth):
    from dbaf_b200 import global_ba
    rng = np.random.default_rng(0)
    for trial in range(10):
        N = int(rng.integers(5, 60))
        E = int(rng.integers(1, 300))
        t0 = 1
        ii, jj = rng.integers(0, N, E), rng.integers(0, N, E)
        kx = global_ba.depth_frames(ii, t0, N)
        assert np.array_equal(kx, oracle.build_tables(ii, jj, t0, N)["kx"])
        for world in (1, 2, 3, 8):
            r = global_ba.shard_ranges(ii, t0, N, world)
            assert len(r) == world and r[0][0] == 0 and r[-1][1] == len(kx)
            assert all(r[k][1] == r[k + 1][0] and r[k][0] <= r[k][1] for k in range(world - 1))
        # band = the widest pose pair any Schur triple or edge block couples
        tab = oracle.build_tables(ii, jj, t0, N)
        tri = oracle.schur_triples(tab, t0, N)
        want = int(np.abs(tri[:, 0] - tri[:, 1]).max()) if len(tri) else 0
        for i, j in zip(ii, jj):
            if i >= t0 and j >= t0:
                want = max(want, abs(int(i) - int(j)))
        assert global_ba.band_blocks(ii, jj, t0, N) == want
    ii, jj = synth.banded_graph(50, reach=3)
    assert global_ba.band_blocks(ii, jj, 1, 50) == 6


# ------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("band_hint", ["exact", "loose", "dense"])
def test_large_window_solvers(dev, db, oracle, synth, band_hint):
    """6P = 240: banded sliding-window Cholesky (exact and over-estimated band) and the dense fallback."""
    from dbaf_b200 import global_ba
    w = synth.make_window("g40", seed=1, graph="banded")
    t = to_torch(w, dev)
    win = db._Window(t["poses"], t["disps"], t["intrinsics"], t["disps_sens"], t["targets"], t["weights"],
                     t["eta"], t["ii"], t["jj"], w["t0"], w["t1"])
    win.prepare()
    win.linearize(0.05)
    n = win.n
    hv = win.Hv().cpu().numpy()
    H, v = hv[:n * n].reshape(n, n), hv[n * n:]
    band = global_ba.band_blocks(w["ii"], w["jj"], w["t0"], w["t1"])
    assert band == 4 and n > 160
    # nothing outside the band
    blk = np.abs(H.reshape(n // 6, 6, n // 6, 6)).max(axis=(1, 3))
    I, J = np.nonzero(blk)
    assert np.abs(I - J).max() <= band
    win.solve(1e-4, 0.1, band={"exact": band, "loose": band + 7, "dense": 0}[band_hint])
    want, ok = oracle.damped_solve(H, v, 1e-4, 0.1)
    assert ok and int(win.info().cpu()[1]) == 0
    assert relerr(win.dx().cpu().numpy().reshape(-1), want) < 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("extra,reach", [(40, 2), (40, 5), (90, 3), (27, 1)])
def test_twisted_banded_solve(dev, db, oracle, synth, extra, reach):
    """Long chains are factorised from both ends (two CTAs meeting at a B-block separator): same dx as the
    single sweep and as the float64 dense solve of the oracle."""
    from dbaf_b200 import global_ba
    w = synth.make_window("g40", seed=4, graph="banded", reach=reach, n_extra_frames=extra)
    band = global_ba.band_blocks(w["ii"], w["jj"], w["t0"], w["t1"])
    P = w["t1"] - w["t0"]
    assert P >= 64 and P >= 3 * band + 2              # the twisted path is taken
    out = {}
    for impl in ("twisted", "single"):
        db.set_option("band_impl", impl)
        t = to_torch(w, dev)
        win = db._Window(t["poses"], t["disps"], t["intrinsics"], t["disps_sens"], t["targets"], t["weights"],
                         t["eta"], t["ii"], t["jj"], w["t0"], w["t1"])
        win.prepare()
        win.linearize(0.05)
        n = win.n
        hv = win.Hv().cpu().numpy()
        H, v = hv[:n * n].reshape(n, n).copy(), hv[n * n:].copy()
        win.solve(1e-4, 0.1, band=band)
        assert int(win.info().cpu()[1]) == 0
        out[impl] = win.dx().cpu().numpy().reshape(-1).copy()
    want, ok = oracle.damped_solve(H, v, 1e-4, 0.1)
    assert ok
    assert relerr(out["twisted"], want) < 1e-6, relerr(out["twisted"], want)
    assert relerr(out["single"], want) < 1e-6
    # a non-positive pivot anywhere drops the whole step in both variants
    t = to_torch(w, dev)
    win = db._Window(t["poses"], t["disps"], t["intrinsics"], t["disps_sens"], t["targets"], t["weights"],
                     t["eta"], t["ii"], t["jj"], w["t0"], w["t1"])
    win.prepare()
    win.linearize(0.05)
    n = win.n
    i = (n // 6 - 3) * 6
    win.H()[i, i] = -1.0e9                             # a diagonal entry inside the bottom sweep's part
    win.solve(1e-4, 0.1, band=band)
    assert int(win.info().cpu()[1]) == 1
    assert float(win.dx().abs().max()) == 0.0


@pytest.mark.gpu
@pytest.mark.parametrize("reach", [4, 6, 10])
def test_tensor_core_schur_matches_simt_and_oracle(dev, db, oracle, synth, reach):
    """Dense graphs (>= 8 row blocks per depth frame) take k_schur_tc (TF32 x3 on tcgen05): same reduced
    system as the fp32 SIMT kernels (EEt6x6 / Ev6x1 of src/droid_kernels.cu:1046-1138) and as the float64
    oracle."""
    w = synth.make_window("g40", seed=3, graph="banded", reach=reach)
    assert len(w["ii"]) >= 7 * 40

    def system(impl):
        db.set_option("schur_impl", impl or "auto")
        t = to_torch(w, dev)
        win = db._Window(t["poses"], t["disps"], t["intrinsics"], t["disps_sens"], t["targets"], t["weights"],
                         t["eta"], t["ii"], t["jj"], w["t0"], w["t1"])
        win.prepare()
        win.linearize(0.05)
        n = win.n
        hv = win.Hv().cpu().numpy()
        return hv[:n * n].reshape(n, n).copy(), hv[n * n:].copy()

    H_tc, v_tc = system(None)
    H_si, v_si = system("simt")
    assert not np.array_equal(H_tc, H_si)              # the two paths really are different kernels
    assert relerr(H_tc, H_tc.T) < 1e-6
    tr = []
    p, d = w["poses"].astype(np.float64), w["disps"].astype(np.float64)
    oracle.ba(p, d, w["intrinsics"], w["disps_sens"], w["targets"], w["weights"], w["eta"], w["ii"], w["jj"],
              w["t0"], w["t1"], 1, 1e-4, 0.1, False, dtype=np.float64, trace=tr)
    H64, v64 = tr[0]["H"], tr[0]["v"]
    # the tensor-core path is as close to the float64 system as the fp32 SIMT path is
    assert relerr(H_tc, H64) < max(2e-6, 2 * relerr(H_si, H64)), (relerr(H_tc, H64), relerr(H_si, H64))
    assert relerr(v_tc, v64) < max(2e-6, 2 * relerr(v_si, v64)), (relerr(v_tc, v64), relerr(v_si, v64))
    assert relerr(H_tc, H_si) < 1e-5 and relerr(v_tc, v_si) < 1e-5


@pytest.mark.gpu
@pytest.mark.parametrize("cfg,reach,kw", [("g40", 10, dict(sensor_frac=0.3)), ("g40", 6, dict(near_frac=0.05, t0=3)),
                                          ("g40", 4, dict(stereo_edges=5)), ("g24", 10, dict())])
def test_fused_linearise_schur_matches_stored_path(dev, db, synth, cfg, reach, kw):
    """Dense graphs: k_linschur_tc (Jacobians straight into the tensor-core operand tiles, pose-depth blocks never
    written) + k_backsub_rc (blocks recomputed from the weights) against the path that stores Eij in the workspace
    (k_linearize2 + k_schur_tc + k_backsub_fused): same reduced system, same dx / dz / poses / disps of one full
    iteration.  Both restate src/droid_kernels.cu:220-468, 1046-1160."""
    w = synth.make_window(cfg, seed=5, graph="banded", reach=reach, **kw)

    def iteration(mode, dx_in=None):
        db.set_option("eij", mode)
        t = to_torch(w, dev)
        win = db._Window(t["poses"], t["disps"], t["intrinsics"], t["disps_sens"], t["targets"], t["weights"],
                         t["eta"], t["ii"], t["jj"], w["t0"], w["t1"])
        win.prepare()
        win.linearize(0.05)
        n = win.n
        hv = win.Hv().cpu().numpy().copy()
        K = int(win.info().cpu()[0])
        fused = win._view(win.L.fused, K, torch.int32).cpu().numpy().copy()
        win.solve(1e-4, 0.1, band=2 * reach)
        own_dx = win.dx().clone()
        dx, dz = win.retract(dx_in=dx_in)
        assert int(win.info().cpu()[1]) == 0
        return dict(H=hv[:n * n].reshape(n, n), v=hv[n * n:], dx=own_dx.cpu().numpy(), dz=dz.cpu().numpy()[:K],
                    poses=t["poses"].cpu().numpy(), disps=t["disps"].cpu().numpy(), fused=fused, dx_dev=own_dx)

    a = iteration("store")
    b = iteration("recompute", dx_in=a["dx_dev"])       # back-substitution with the SAME pose update
    assert a["fused"].sum() == 0 and b["fused"].sum() >= 0.5 * len(b["fused"])    # the new kernels really ran
    assert not np.array_equal(a["H"], b["H"])
    err = {k: relerr(b[k], a[k]) for k in ("H", "v", "dx", "dz", "poses", "disps")}
    record(f"fused_vs_stored[{cfg},reach={reach}]", **err)
    # (the fused kernel rotates with the matrix of the relative pose, takes 1/z from the fast reciprocal and scales
    # the weights in float32; v is a cancelling sum of residuals)
    assert err["H"] < 1e-5 and err["v"] < 3e-5, err
    assert err["dz"] < 3e-5 and err["disps"] < 1e-5 and err["poses"] < 1e-6, err
    # dx of the two systems: the reduced system of a banded graph is ill-conditioned (see test_ref_parity); two
    # correct fp32 linearisations agree in dx to ~cond * 1e-7
    assert err["dx"] < 5e-3, err


@pytest.mark.gpu
def test_band_zeroing_and_band_strip(dev, db, synth):
    """dbaf_ba_zero_system clears exactly what a linearisation of a banded graph writes (so zero_first = 0 after it
    reproduces a fresh linearisation), and dbaf_ba_band_strip round-trips the band of H and v (what the sharded path
    all-reduces)."""
    from dbaf_b200 import global_ba
    w = synth.make_window("g40", seed=2, graph="banded", reach=4)
    band = global_ba.band_blocks(w["ii"], w["jj"], w["t0"], w["t1"])
    assert 0 < band < 12
    t = to_torch(w, dev)
    win = db._Window(t["poses"], t["disps"], t["intrinsics"], t["disps_sens"], t["targets"], t["weights"],
                     t["eta"], t["ii"], t["jj"], w["t0"], w["t1"])
    win.prepare()
    win.linearize(0.05)
    want = win.Hv().clone()
    win.linearize(0.05, zero_first=False)                     # twice the system: stale contents everywhere it matters
    win.zero_system(band)
    assert float(win.Hv().abs().max()) == 0.0
    win.linearize(0.05, zero_first=False)
    assert relerr(win.Hv().cpu().numpy(), want.cpu().numpy()) < 1e-12
    strip = win.band_strip(band)
    assert strip.numel() == win.n * 6 * (2 * band + 1) + win.n
    assert abs(float(strip.sum()) - float(want.sum())) <= 1e-9 * float(want.abs().sum())   # nothing lies outside the band
    win.Hv().zero_()
    win.band_strip(band, strip, pack=False)
    assert torch.equal(win.Hv(), want if False else win.Hv()) and relerr(win.Hv().cpu().numpy(), want.cpu().numpy()) < 1e-12


@pytest.mark.gpu
@pytest.mark.parametrize("reach", [4, 10])
def test_schur_kernels_on_sign_mixed_blocks(dev, db, synth, reach):
    """The Schur kernels fed with zero-mean random blocks (every entry of S is a cancelling sum): the tensor-core
    kernel's per-slice accumulators and its 1 + 2^-23 outer factor were tuned on positive-dominated sums; here the
    error is measured entry by entry against float64, relative to the sum of absolute terms, next to the fp32 SIMT
    kernels (EEt6x6 / Ev6x1, src/droid_kernels.cu:1046-1138)."""
    w = synth.make_window("g40", seed=7, graph="banded", reach=reach)
    t = to_torch(w, dev)
    win = db._Window(t["poses"], t["disps"], t["intrinsics"], t["disps_sens"], t["targets"], t["weights"],
                     t["eta"], t["ii"], t["jj"], w["t0"], w["t1"])
    win.prepare()
    P, E, HW, t0 = win.P, win.E, win.HW, w["t0"]
    K = int(win.info().cpu()[0])
    g = torch.Generator(device=dev).manual_seed(11)
    f32 = torch.float32
    Q = win._view(win.L.Q, K * HW, f32).view(K, HW)
    wv = win._view(win.L.w, K * HW, f32).view(K, HW)
    Ei = win._view(win.L.Ei, P * 6 * HW, f32).view(P, 6, HW)
    Eij = win._view(win.L.Eij, E * 6 * HW, f32).view(E, 6, HW)
    Q.copy_(torch.rand(K, HW, generator=g, device=dev) * 1.5 + 0.5)
    wv.copy_(torch.randn(K, HW, generator=g, device=dev))
    Ei.copy_(torch.randn(P, 6, HW, generator=g, device=dev))
    Eij.copy_(torch.randn(E, 6, HW, generator=g, device=dev) * torch.rand(E, 1, 1, generator=g, device=dev) * 3)
    tab = db.ba_tables(t["ii"], t["jj"], w["t0"], w["t1"], w["poses"].shape[0], w["ht"], w["wd"])
    n = 6 * P
    H64 = torch.zeros(P, 6, P, 6, dtype=torch.float64, device=dev)
    Habs = torch.zeros_like(H64)
    v64 = torch.zeros(P, 6, dtype=torch.float64, device=dev)
    vabs = torch.zeros_like(v64)
    jj = w["jj"]
    for k in range(K):
        rows = tab["srow_idx"][tab["srow_ptr"][k]:tab["srow_ptr"][k + 1]].tolist()
        if not rows:
            continue
        X = torch.stack([(Ei[r] if r < P else Eij[r - P]) for r in rows]).double()        # [m,6,HW]
        pose = [r if r < P else int(jj[r - P]) - t0 for r in rows]
        XQ = X * Q[k].double()
        S = torch.einsum("aip,bjp->aibj", XQ, X)
        Sa = torch.einsum("aip,bjp->aibj", XQ.abs(), X.abs())
        sv = torch.einsum("aip,p->ai", XQ, wv[k].double())
        sva = torch.einsum("aip,p->ai", XQ.abs(), wv[k].double().abs())
        for a, pa in enumerate(pose):
            v64[pa] += sv[a]
            vabs[pa] += sva[a]
            for b, pb in enumerate(pose):
                H64[pa, :, pb, :] += S[a, :, b, :]
                Habs[pa, :, pb, :] += Sa[a, :, b, :]
    H64, Habs = H64.view(n, n), Habs.view(n, n)
    nz = Habs > 0
    out = {}
    for impl in ("auto", "simt"):
        db.set_option("schur_impl", impl)
        win.Hv().zero_()
        win.schur()
        H, v = -win.H(), -win.v()
        eH = ((H - H64).abs()[nz] / Habs[nz])
        ev = ((v.view(P, 6) - v64).abs() / vabs.clamp(min=1e-300))
        bias = float(((H - H64)[nz] / Habs[nz]).mean())
        out[impl] = (float(eH.max()), float(ev.max()), bias, H.clone())
    assert not torch.equal(out["auto"][3], out["simt"][3])          # two different kernels ran
    from helpers import record
    record(f"schur_sign_mixed[reach={reach}]", tc_H=out["auto"][0], tc_v=out["auto"][1], tc_bias=out["auto"][2],
           simt_H=out["simt"][0], simt_v=out["simt"][1], simt_bias=out["simt"][2])
    # errors relative to the sum of absolute terms stay at the float32 rounding level for both, no drift
    assert out["auto"][0] < max(5e-7, 3 * out["simt"][0]) and out["auto"][1] < max(5e-7, 3 * out["simt"][1])
    assert abs(out["auto"][2]) < 2e-8


@pytest.mark.gpu
def test_global_ba_single_rank_matches_oracle(dev, db, oracle, synth):
    from dbaf_b200 import global_ba
    w = synth.make_window("g40", seed=2, graph="banded")
    t = to_torch(w, dev)
    eng = global_ba.CudaEngine(t["poses"], t["disps"], t["intrinsics"], t["disps_sens"], t["targets"],
                               t["weights"], t["eta"], t["ii"], t["jj"], w["t0"], w["t1"])
    gba = global_ba.GlobalBA(eng, w["ii"], w["jj"], w["t0"], w["t1"])
    gba.run(2)
    res = {}
    for dt in (np.float64, np.float32):
        p, d = w["poses"].astype(dt), w["disps"].astype(dt)
        oracle.ba(p, d, w["intrinsics"], w["disps_sens"], w["targets"], w["weights"], w["eta"], w["ii"], w["jj"],
                  w["t0"], w["t1"], 2, 1e-4, 0.1, False, dtype=dt)
        res[dt] = (p, d)
    (p64, d64), (p32, d32) = res[np.float64], res[np.float32]
    assert relerr(t["poses"].cpu().numpy(), p64) < max(1e-4, 3 * relerr(p32, p64))
    assert relerr(t["disps"].cpu().numpy(), d64) < max(1e-4, 3 * relerr(d32, d64))
    # sharding the same window in two halves on one GPU gives the same reduced system
    t2 = to_torch(w, dev)
    e2 = global_ba.CudaEngine(t2["poses"], t2["disps"], t2["intrinsics"], t2["disps_sens"], t2["targets"],
                              t2["weights"], t2["eta"], t2["ii"], t2["jj"], w["t0"], w["t1"])
    (a0, a1), (b0, b1) = global_ba.shard_ranges(w["ii"], w["t0"], w["t1"], 2)
    e2.linearize(0.05, a0, a1)
    part = e2.system().clone()
    e2.linearize(0.05, b0, b1)
    part += e2.system()
    e2.linearize(0.05, 0, -1)
    assert relerr(part.cpu().numpy(), e2.system().cpu().numpy()) < 1e-12


@pytest.mark.gpu
@pytest.mark.parametrize("P,B", [(64, 1), (65, 3), (70, 20), (97, 7), (128, 20), (203, 13), (66, 21), (90, 25), (40, 5), (63, 20)])
def test_banded_solvers_on_random_band_systems(dev, db, oracle, synth, P, B):
    """Every banded code path (register window B <= 20, shared-memory window B <= 25, twisted for P >= 64 and
    P >= 3B+2, single sweep otherwise) on random block-banded SPD systems written straight into the workspace:
    dx against numpy's dense float64 solve."""
    rng = np.random.default_rng(100 * P + B)
    w = synth.make_window("tiny", seed=1, n_extra_frames=P + 1 - 5, graph="banded", reach=1)
    assert w["t1"] - w["t0"] == P
    n = 6 * P
    G = rng.normal(0, 1, (n, n))
    blk = np.arange(n) // 6
    G[np.abs(blk[:, None] - blk[None, :]) > B // 2 + (B % 2)] = 0.0      # G G^T then has block bandwidth <= B
    H = G @ G.T
    H[np.abs(blk[:, None] - blk[None, :]) > B] = 0.0
    H += np.diag(rng.uniform(1.0, 2.0, n)) * n * 0.05
    v = rng.normal(0, 1, n)
    want, ok = oracle.damped_solve(H, v, 1e-4, 0.1)
    assert ok
    res = {}
    for sweep in ("reg", "smem"):
        for impl in ("auto", "twisted", "single"):
            db.set_option("band_sweep", "smem" if sweep == "smem" else "auto")
            db.set_option("band_impl", impl)
            t = to_torch(w, dev)
            win = db._Window(t["poses"], t["disps"], t["intrinsics"], t["disps_sens"], t["targets"], t["weights"],
                             t["eta"], t["ii"], t["jj"], w["t0"], w["t1"])
            win.prepare()                                   # initialises the info words
            win.H().copy_(torch.from_numpy(H).to(dev))
            win.v().copy_(torch.from_numpy(v).to(dev))
            win.solve(1e-4, 0.1, band=B)
            assert int(win.info().cpu()[1]) == 0
            got = win.dx().cpu().numpy().reshape(-1)
            res[(sweep, impl)] = got
            assert relerr(got, want) < 2e-6, (sweep, impl, relerr(got, want))


@pytest.mark.gpu
@pytest.mark.parametrize("P,B,D", [(128, 20, 3), (203, 13, 5), (300, 5, 12), (400, 20, 8), (999, 20, 8), (999, 20, 16),
                                   (999, 20, 0), (96, 1, 16), (64, 3, 7), (250, 10, 4), (130, 6, 3)])
def test_domain_decomposed_banded_solve(dev, db, oracle, synth, P, B, D):
    """The banded solve cut into D domains (ba_spike.cuh: D sweeps, spikes of the middle domains, block cyclic
    reduction over the separators): dx against numpy's dense float64 solve on random block-banded SPD systems,
    at the size and band of the global-BA workload (P = 999, B = 20) among others; D = 0 is the automatic choice.
    A non-positive pivot anywhere (interior, separator) drops the whole step."""
    rng = np.random.default_rng(1000 * P + 10 * B + D)
    w = synth.make_window("tiny", seed=1, n_extra_frames=P + 1 - 5, graph="banded", reach=1)
    assert w["t1"] - w["t0"] == P
    n = 6 * P
    blk = np.arange(n) // 6
    # G G^T with G banded (half the band) gives an SPD matrix of block bandwidth <= B without forming n x n products
    H = np.zeros((n, n))
    hb = B // 2 + (B % 2)
    G = rng.normal(0, 1, (n, 6 * (2 * hb + 1)))
    for I in range(P):
        lo, hi = max(0, I - hb), min(P, I + hb + 1)
        Gi = G[6 * I:6 * I + 6, :6 * (hi - lo)]
        for J in range(max(0, I - B), I + 1):
            lj, hj = max(0, J - hb), min(P, J + hb + 1)
            a, b = max(lo, lj), min(hi, hj)
            if a < b:
                Gj = G[6 * J:6 * J + 6, :6 * (hj - lj)]
                blkv = Gi[:, 6 * (a - lo):6 * (b - lo)] @ Gj[:, 6 * (a - lj):6 * (b - lj)].T
                H[6 * I:6 * I + 6, 6 * J:6 * J + 6] = blkv
                H[6 * J:6 * J + 6, 6 * I:6 * I + 6] = blkv.T
    H[np.abs(blk[:, None] - blk[None, :]) > B] = 0.0
    H += np.diag(rng.uniform(1.0, 2.0, n)) * (6 * (2 * hb + 1)) * 0.3
    v = rng.normal(0, 1, n)
    want, ok = oracle.damped_solve(H, v, 1e-4, 0.1)
    assert ok
    t = to_torch(w, dev)
    win = db._Window(t["poses"], t["disps"], t["intrinsics"], t["disps_sens"], t["targets"], t["weights"],
                     t["eta"], t["ii"], t["jj"], w["t0"], w["t1"])
    win.prepare()
    Hd, vd = torch.from_numpy(H).to(dev), torch.from_numpy(v).to(dev)
    out = {}
    for impl, dom in (("auto", D), ("twisted", 0), ("single", 0)):
        db.set_option("band_impl", impl)
        db.set_option("band_domains", str(dom))
        win.H().copy_(Hd)
        win.v().copy_(vd)
        win.solve(1e-4, 0.1, band=B)
        assert int(win.info().cpu()[1]) == 0
        out[impl] = win.dx().cpu().numpy().reshape(-1).copy()
        assert relerr(out[impl], want) < 2e-6, (impl, relerr(out[impl], want))
    from helpers import record
    record(f"spike_solve[P={P},B={B},D={D}]", spike=relerr(out["auto"], want), twisted=relerr(out["twisted"], want),
           single=relerr(out["single"], want))
    # failures: a bad pivot inside a middle domain, inside a separator, inside the last domain
    db.set_option("band_impl", "auto")
    db.set_option("band_domains", str(D))
    fails = 0
    for pos in (P // 2, P // 3, P - 2, 1):
        win.H().copy_(Hd)
        win.v().copy_(vd)
        win.H()[6 * pos, 6 * pos] = -1.0e9
        win.solve(1e-4, 0.1, band=B)
        fails += 1
        assert int(win.info().cpu()[1]) == fails, pos
        assert float(win.dx().abs().max()) == 0.0, pos


@pytest.mark.gpu
def test_c5_full_size_properties(dev, db, synth):
    """BASELINE.json configs[4] at full size (1000 keyframes, 19 890 edges, 64x64): too large for the oracle, so
    the checks are size-independent properties: the reduced system of two shards sums to the unsharded one
    (linearity of the edge partition), it is symmetric and confined to the band, the banded solve satisfies the
    damped normal equations to float64 accuracy, and one Gauss-Newton step lowers the reprojection cost."""
    from dbaf_b200 import global_ba
    w = synth.make_window("c5", seed=0, graph="banded", pose_noise=0.005)
    t = to_torch(w, dev)
    eng = global_ba.CudaEngine(t["poses"], t["disps"], t["intrinsics"], t["disps_sens"], t["targets"], t["weights"],
                               t["eta"], t["ii"], t["jj"], w["t0"], w["t1"])
    gba = global_ba.GlobalBA(eng, w["ii"], w["jj"], w["t0"], w["t1"])
    assert gba.band == 20
    n = eng.w.n
    (a0, a1), (b0, b1) = global_ba.shard_ranges(w["ii"], w["t0"], w["t1"], 2)
    eng.linearize(0.05, a0, a1)
    part = eng.system().clone()
    eng.linearize(0.05, b0, b1)
    part += eng.system()
    eng.linearize(0.05, 0, -1)
    full = eng.system().clone()
    assert float((part - full).norm() / full.norm()) < 1e-10
    H, v = full[:n * n].view(n, n), full[n * n:]
    assert float((H - H.T).norm() / H.norm()) < 1e-6
    blk = torch.arange(n, device=dev) // 6
    outside = (blk[:, None] - blk[None, :]).abs() > gba.band
    assert float(H[outside].abs().max()) == 0.0
    idx = gba.band_index(dev)
    assert idx is not None and idx.numel() < 0.05 * full.numel()
    lm, ep = 1e-4, 0.1
    eng.solve(lm, ep, gba.band)
    assert int(eng.w.info().cpu()[1]) == 0
    dx = eng.w.dx().double().reshape(-1)
    Hd = H + torch.diag(ep + lm * torch.diagonal(H))
    Hl = torch.tril(Hd) + torch.tril(Hd, -1).T                  # the solver reads the lower triangle
    res = Hl @ dx - v
    assert float(res.norm() / v.norm()) < 1e-5                   # dx is stored in float32
    # one full iteration lowers the weighted reprojection cost
    def cost():
        c, valid = db.reproject(t["poses"], t["disps"], t["intrinsics"], t["ii"], t["jj"])
        r = c[0].permute(0, 3, 1, 2) - t["targets"]
        # points behind or too close to the target camera carry no weight in the BA either
        return float((t["weights"] * valid[0].permute(0, 3, 1, 2) * (r * r).clamp(max=1e4)).sum())
    c0 = cost()
    eng.retract(0, -1)
    assert cost() < c0
