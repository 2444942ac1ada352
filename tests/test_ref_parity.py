"""Differential tests against the UNMODIFIED reference CUDA extension (oracle/_ref, built from
/root/reference/src by oracle/build_ref.py) on the same GPU and the same inputs.  Skipped when the
extension was never built.  This is the strongest parity statement the repo makes: the reference
has no tests or golden vectors of its own."""
import numpy as np
import pytest
import torch

from helpers import ba_args, record, relerr, to_torch

pytestmark = pytest.mark.gpu
TOL = 1e-4


def _np(t):
    return t.detach().cpu().numpy()


@pytest.mark.parametrize("cfg,kw", [
    ("tiny", dict(seed=1, sensor_frac=0.3, near_frac=0.05, stereo_edges=2)),
    ("c2", dict(seed=0)),
    ("c2_real", dict(seed=1, sensor_frac=0.3)),
    ("c3", dict(seed=2, near_frac=0.02)),
])
def test_bacore_hessian_and_retract(dev, db, ref, synth, cfg, kw):
    w = synth.make_window(cfg, **kw)
    n = 6 * (w["t1"] - w["t0"])
    outs = []
    for mod in (db, ref):
        t = to_torch(w, dev)
        core = mod.BACore()
        core.init(*ba_args(t), w["t0"], w["t1"], 2, 1e-4, 0.1, False)
        H = torch.zeros(n, n, dtype=torch.float64)
        v = torch.zeros(n, dtype=torch.float64)
        core.hessian(H, v)
        dx = torch.from_numpy(np.random.default_rng(5).normal(0, 1e-3, n))
        dxr, dzr = core.retract(dx)
        outs.append((H.numpy().copy(), v.numpy().copy(), _np(dxr), _np(dzr), _np(t["poses"]), _np(t["disps"])))
        del core
    (H, v, dx, dz, p, d), (Hr, vr, dxr, dzr, pr, dr) = outs
    record(f"bacore_vs_reference[{cfg}]", H=relerr(H, Hr), v=relerr(v, vr), dz_given_dx=relerr(dz, dzr),
           poses=relerr(p, pr), disps=relerr(d, dr))
    assert relerr(H, Hr) < TOL and relerr(v, vr) < TOL
    assert relerr(dx, dxr) < 1e-6 and relerr(dz, dzr) < TOL
    assert relerr(p, pr) < 1e-6 and relerr(d, dr) < TOL


@pytest.mark.parametrize("cfg,kw", [("tiny", dict(seed=2)), ("c2", dict(seed=0)), ("c3", dict(seed=1))])
@pytest.mark.parametrize("motion_only", [False, True])
def test_ba(dev, db, ref, oracle, synth, cfg, kw, motion_only):
    w = synth.make_window(cfg, **kw)
    if motion_only:
        # The reference aborts the process (Eigen assertion in SparseMatrix::diagonal(), reached from
        # SparseBlock::solve, src/droid_kernels.cu:1253) when a window pose has no edge at all in
        # motion-only mode; ours returns dx = 0 for such a pose (covered in test_gpu_parity).
        touched = set(w["ii"].tolist()) | set(w["jj"].tolist())
        if not all(f in touched for f in range(w["t0"], w["t1"])):
            pytest.skip("window has an isolated pose: the reference aborts on it in motion-only mode")
    res = []
    for mod in (db, ref):
        t = to_torch(w, dev)
        dx, dz = mod.ba(*ba_args(t), w["t0"], w["t1"], 2, 1e-4, 0.1, motion_only)
        res.append((_np(dx), None if dz is None else _np(dz), _np(t["poses"]), _np(t["disps"])))
    (dx, dz, p, d), (dxr, dzr, pr, dr) = res
    # float64 answer to measure how far the reference itself is from the exact solution
    p64, d64 = w["poses"].astype(np.float64), w["disps"].astype(np.float64)
    dx64, dz64 = oracle.ba(p64, d64, w["intrinsics"], w["disps_sens"], w["targets"], w["weights"], w["eta"],
                           w["ii"], w["jj"], w["t0"], w["t1"], 2, 1e-4, 0.1, motion_only, dtype=np.float64)
    slack = lambda r, x: max(TOL, 3 * relerr(r, x))      # see test_gpu_parity.test_ba_vs_oracle
    record(f"ba_vs_reference[{cfg},motion_only={motion_only}]", dx=relerr(dx, dxr), poses=relerr(p, pr),
           dz=None if motion_only else relerr(dz, dzr), disps=None if motion_only else relerr(d, dr),
           ref_dx_vs_f64=relerr(dxr, dx64), ours_dx_vs_f64=relerr(dx, dx64),
           ref_dz_vs_f64=None if motion_only else relerr(dzr, dz64),
           ours_dz_vs_f64=None if motion_only else relerr(dz, dz64))
    assert relerr(dx, dxr) < slack(dxr, dx64), (relerr(dx, dxr), relerr(dxr, dx64), relerr(dx, dx64))
    assert relerr(p, pr) < slack(pr, p64)
    if motion_only:
        assert dz is None and dzr is None
    else:
        assert relerr(dz, dzr) < slack(dzr, dz64), (relerr(dz, dzr), relerr(dzr, dz64))
        assert relerr(d, dr) < slack(dr, d64)


@pytest.mark.parametrize("cfg,kw,prior_n,iters", [("tiny", dict(seed=5), 12, 1), ("tiny", dict(seed=6, sensor_frac=0.3), None, 2),
                                                  ("c2", dict(seed=0), 18, 2)])
def test_ba_extend_vs_reference(dev, db, ref, oracle, synth, cfg, kw, prior_n, iters):
    """ba_extend against ba_cuda_extend itself (src/droid_kernels.cu:1514-1678): the exported undamped H, v, the
    prior arithmetic of :1618-1644 (L = Ad + Adprior with Adprior = Ad overwritten by A_prior on its leading block),
    the resulting updates, and skip_solve."""
    w = synth.make_window(cfg, **kw)
    n = 6 * (w["t1"] - w["t0"])
    rng = np.random.default_rng(6)
    pn = n if prior_n is None else prior_n
    M = rng.normal(0, 1, (pn, pn))
    # a prior that dominates the block it replaces, as a marginalisation prior does: L = Ad + Adprior doubles every
    # block of A - S outside the prior's, so a weak prior leaves L indefinite and both sides return dx = 0
    t = to_torch(w, dev)
    H0 = torch.zeros(n, n, dtype=torch.float64)
    db.ba_extend(*ba_args(t), H0, torch.zeros(n, dtype=torch.float64), torch.zeros(0, 0, dtype=torch.float64),
                 w["t0"], w["t1"], 1, 1e-4, 0.1, False, True)
    prior = torch.from_numpy((M @ M.T / pn + np.eye(pn)) * 4.0 * float(H0.diagonal().max()))
    res = []
    for mod in (db, ref):
        t = to_torch(w, dev)
        H = torch.zeros(n, n, dtype=torch.float64)
        v = torch.zeros(n, dtype=torch.float64)
        dx, dz = mod.ba_extend(*ba_args(t), H, v, prior, w["t0"], w["t1"], iters, 1e-4, 0.1, False, False)
        res.append((H.numpy().copy(), v.numpy().copy(), _np(dx), _np(dz), _np(t["poses"]), _np(t["disps"])))
    (H, v, dx, dz, p, d), (Hr, vr, dxr, dzr, pr, dr) = res
    # yardstick for dx: the same reduced system solved in float64 from either linearisation
    def f64_solve(Hm, vm):
        A = Hm.copy()
        A[:pn, :pn] += prior.numpy()
        mask = np.ones_like(A, bool)
        mask[:pn, :pn] = False
        A[mask] *= 2.0
        return oracle.damped_solve(A, vm, 1e-4, 0.1)[0]
    yard = relerr(f64_solve(H, v), f64_solve(Hr, vr))
    record(f"ba_extend_vs_reference[{cfg},prior={pn},iters={iters}]", H=relerr(H, Hr), v=relerr(v, vr),
           dx=relerr(dx, dxr), dz=relerr(dz, dzr), poses=relerr(p, pr), disps=relerr(d, dr), dx_f64_yardstick=yard)
    assert relerr(H, Hr) < TOL and relerr(v, vr) < TOL
    assert np.abs(dxr).max() > 0 and np.abs(dx).max() > 0          # the factorisation succeeded on both sides
    slack = max(TOL, 3 * yard)
    assert relerr(dx, dxr) < slack and relerr(p, pr) < slack
    assert relerr(dz, dzr) < slack and relerr(d, dr) < slack
    if iters == 1:
        # given the exported system, the prior arithmetic and the float64 solve are exact to 1e-5
        assert relerr(dx.reshape(-1), f64_solve(H, v)) < 1e-5
    # skip_solve: export only, nothing updated, [None, None] back
    first = []
    for mod in (db, ref):
        t = to_torch(w, dev)
        H2 = torch.zeros(n, n, dtype=torch.float64)
        v2 = torch.zeros(n, dtype=torch.float64)
        out = mod.ba_extend(*ba_args(t), H2, v2, prior, w["t0"], w["t1"], 1, 1e-4, 0.1, False, True)
        assert list(out) == [None, None]
        assert torch.equal(t["poses"].cpu(), torch.from_numpy(w["poses"])) and torch.equal(t["disps"].cpu(), torch.from_numpy(w["disps"]))
        first.append((H2.numpy(), v2.numpy()))
    assert relerr(first[0][0], first[1][0]) < TOL and relerr(first[0][1], first[1][1]) < TOL
    # an empty prior leaves Adprior = Ad, i.e. the reference solves with 2 (A - S) (:1622-1644)
    res0 = []
    for mod in (db, ref):
        t = to_torch(w, dev)
        H2 = torch.zeros(n, n, dtype=torch.float64)
        v2 = torch.zeros(n, dtype=torch.float64)
        dx0, _ = mod.ba_extend(*ba_args(t), H2, v2, torch.zeros(0, 0, dtype=torch.float64), w["t0"], w["t1"], 1, 1e-4,
                               0.1, False, False)
        res0.append(_np(dx0))
    assert relerr(res0[0], res0[1]) < slack
    # a smaller H / v receives the leading block only (:1632-1640)
    t = to_torch(w, dev)
    Hs = torch.zeros(12, 12, dtype=torch.float64)
    vs = torch.zeros(12, dtype=torch.float64)
    db.ba_extend(*ba_args(t), Hs, vs, prior, w["t0"], w["t1"], 1, 1e-4, 0.1, False, True)
    t = to_torch(w, dev)
    Hsr = torch.zeros(12, 12, dtype=torch.float64)
    vsr = torch.zeros(12, dtype=torch.float64)
    ref.ba_extend(*ba_args(t), Hsr, vsr, prior, w["t0"], w["t1"], 1, 1e-4, 0.1, False, True)
    assert relerr(Hs.numpy(), Hsr.numpy()) < TOL and relerr(vs.numpy(), vsr.numpy()) < TOL


def test_c5_matches_reference(dev, db, ref, synth):
    """BASELINE.json configs[4] at full size (1000 keyframes, 19 890 edges, 64x64) against the reference itself:
    the exact kernels the multi-GPU numbers come from (k_linearize2<2,*>, k_schur_tc with one CTA per frame, the
    banded / domain-decomposed solve at B = 20, k_band_back, k_backsub_fused) versus BACore::hessian
    (src/droid_kernels.cu:1816-1900), ba_cuda_extend's export (:1514-1640) and ba_cuda (:1394-1512)."""
    from dbaf_b200 import global_ba
    w = synth.make_window("c5", seed=0, graph="banded", pose_noise=0.005)
    t0, t1 = w["t0"], w["t1"]
    P = t1 - t0
    n = 6 * P
    blk = torch.arange(n, device=dev) // 6

    # ---- (1) BACore.hessian: undamped reduced camera system, alpha = 0.001
    t = to_torch(w, dev)
    core = ref.BACore()
    core.init(*ba_args(t), t0, t1, 1, 1e-4, 0.1, False)
    Hr = torch.zeros(n, n, dtype=torch.float64)
    vr = torch.zeros(n, dtype=torch.float64)
    core.hessian(Hr, vr)
    del core
    eng = global_ba.CudaEngine(*ba_args(t), t0, t1)
    gba = global_ba.GlobalBA(eng, w["ii"], w["jj"], t0, t1)
    assert gba.band == 20
    eng.linearize(0.001, 0, -1)
    sys_ = eng.system()
    H, v = sys_[:n * n].view(n, n), sys_[n * n:]
    Hr_d, vr_d = Hr.to(dev), vr.to(dev)
    outside = (blk[:, None] - blk[None, :]).abs() > gba.band
    assert float(Hr_d[outside].abs().max()) == 0.0 and float(H[outside].abs().max()) == 0.0
    eH = float((H - Hr_d).abs().max() / Hr_d.abs().max())
    ev = float((v - vr_d).abs().max() / vr_d.abs().max())
    assert eH < TOL and ev < TOL, (eH, ev)

    # ---- (2) the system `ba` solves (alpha = 0.05), exported by ba_extend(skip_solve)
    t = to_torch(w, dev)
    Hr5 = torch.zeros(n, n, dtype=torch.float64)
    vr5 = torch.zeros(n, dtype=torch.float64)
    ref.ba_extend(*ba_args(t), Hr5, vr5, torch.zeros(0, 0, dtype=torch.float64), t0, t1, 1, 1e-4, 0.1, False, True)
    eng.linearize(0.05, 0, -1)
    H5, v5 = H.clone(), v.clone()
    Hr5_d, vr5_d = Hr5.to(dev), vr5.to(dev)
    eH5 = float((H5 - Hr5_d).abs().max() / Hr5_d.abs().max())
    ev5 = float((v5 - vr5_d).abs().max() / vr5_d.abs().max())
    assert eH5 < TOL and ev5 < TOL, (eH5, ev5)

    def f64_solve(Hm, vm):
        A = torch.tril(Hm) + torch.tril(Hm, -1).T
        A = A + torch.diag(0.1 + 1e-4 * torch.diagonal(A))
        return torch.linalg.solve(A, vm)
    dx64_ours, dx64_ref = f64_solve(H5, v5), f64_solve(Hr5_d, vr5_d)
    yard = float((dx64_ours - dx64_ref).abs().max() / dx64_ref.abs().max())

    # ---- (3) one iteration of ba
    tr = to_torch(w, dev)
    dxr, dzr = ref.ba(*ba_args(tr), t0, t1, 1, 1e-4, 0.1, False)
    to = to_torch(w, dev)
    eng = global_ba.CudaEngine(*ba_args(to), t0, t1)
    gba = global_ba.GlobalBA(eng, w["ii"], w["jj"], t0, t1)
    gba.run(1)
    assert int(eng.w.info().cpu()[1]) == 0
    dx = eng.w.dx()
    rel = lambda a, b: float((a.double() - b.double()).abs().max() / b.double().abs().max())
    e_dx, e_p, e_d = rel(dx, dxr), rel(to["poses"], tr["poses"]), rel(to["disps"], tr["disps"])
    # the solver itself: our float32 dx against the float64 solve of our own system
    e_solver = float((dx.double().reshape(-1) - dx64_ours).abs().max() / dx64_ours.abs().max())
    e_ref_solver = float((dxr.double().reshape(-1) - dx64_ref).abs().max() / dx64_ref.abs().max())
    # dz given the same dx: feed the reference's dx through our back-substitution
    t3 = to_torch(w, dev)
    eng3 = global_ba.CudaEngine(*ba_args(t3), t0, t1)
    eng3.linearize(0.05, 0, -1)
    _, dz3 = eng3.w.retract(dx_in=dxr.contiguous(), want_dx=False, want_dz=True)
    e_dz = rel(dz3, dzr)
    record("c5_vs_reference", H_bacore=eH, v_bacore=ev, H_ba=eH5, v_ba=ev5, dx=e_dx, poses=e_p, disps=e_d,
           dz_given_dx=e_dz, dx_f64_yardstick=yard, ours_dx_vs_own_f64=e_solver, ref_dx_vs_own_f64=e_ref_solver)
    assert e_solver < 1e-5 and e_dz < TOL
    slack = max(TOL, 3 * yard)
    assert e_dx < slack and e_p < slack and e_d < slack, (e_dx, e_p, e_d, yard)


def test_geometry(dev, db, ref, synth):
    w = synth.make_window("c2_real", seed=3, near_frac=0.05)
    t = to_torch(w, dev)
    a, b = db.projmap(t["poses"], t["disps"], t["intrinsics"], t["ii"], t["jj"]), \
        ref.projmap(t["poses"], t["disps"], t["intrinsics"], t["ii"], t["jj"])
    assert (a[1] != b[1]).float().mean().item() < 1e-5
    assert relerr(_np(a[0]), _np(b[0])) < 1e-5
    for beta in (0.3, 0.0, 1.0):
        d1 = db.frame_distance(t["poses"], t["disps"], t["intrinsics"], t["ii"], t["jj"], beta)
        d2 = ref.frame_distance(t["poses"], t["disps"], t["intrinsics"], t["ii"], t["jj"], beta)
        assert torch.equal(d1 > 999, d2 > 999)
        assert relerr(_np(d1), _np(d2)) < 1e-4
    assert relerr(_np(db.iproj(t["poses"], t["disps"], t["intrinsics"])),
                  _np(ref.iproj(t["poses"], t["disps"], t["intrinsics"]))) < 1e-6
    tg = to_torch(dict(p=w["poses_gt"], d=w["disps_gt"]), dev)
    ix = torch.arange(w["poses"].shape[0], device=dev)
    th = torch.full((len(ix),), 0.05, device=dev)
    c1 = db.depth_filter(tg["p"], tg["d"], t["intrinsics"], ix, th)
    c2 = ref.depth_filter(tg["p"], tg["d"], t["intrinsics"], ix, th)
    assert (c1 != c2).float().mean().item() < 1e-4


@pytest.mark.parametrize("cfg,n", [("tiny", 4), ("c2", 3), ("c2_real", 3), ("c3", 3)])
def test_corr_lookup_bit_exact(dev, db, ref, synth, cfg, n):
    """fp16 lookups must reproduce the reference bit for bit on every pyramid level."""
    w = synth.make_window(cfg, seed=0)
    f = torch.from_numpy(synth.make_features(cfg, seed=0, n_frames=w["poses"].shape[0])).to(dev)
    coords = torch.from_numpy(synth.lookup_coords(w, 0.05, 0)[:n]).to(dev)
    f1, f2 = f[w["ii"][:n]], f[w["jj"][:n]]
    nn, C, ht, wd = f1.shape
    corr = torch.matmul((f1.reshape(nn, C, -1) / 4.0).transpose(1, 2), f2.reshape(nn, C, -1) / 4.0)
    corr = corr.reshape(nn * ht * wd, 1, ht, wd)
    for l in range(4):
        vol = corr.view(nn, ht, wd, ht >> l, wd >> l).contiguous()
        c = (coords / 2 ** l).contiguous()
        got, = db.corr_index_forward(vol, c, 3)
        want, = ref.corr_index_forward(vol, c, 3)
        assert torch.equal(got.view(torch.int16), want.view(torch.int16)), (cfg, l)
        g32, = db.corr_index_forward(vol.float(), c, 3)
        w32, = ref.corr_index_forward(vol.float(), c, 3)
        assert relerr(_np(g32), _np(w32)) < 1e-6
        corr = torch.nn.functional.avg_pool2d(corr, 2, stride=2)


def test_altcorr(dev, db, ref, synth):
    w = synth.make_window("c4", seed=0)
    f = torch.from_numpy(synth.make_features("c4", seed=0, n_frames=w["poses"].shape[0])).to(dev)
    n = 4
    coords = torch.from_numpy(synth.lookup_coords(w, 0.05, 0)[:n]).to(dev).permute(0, 2, 3, 1)[:, None].contiguous()
    f1 = (f[w["ii"][:n]].float() / 4.0).permute(0, 2, 3, 1).contiguous()
    f2 = (f[w["jj"][:n]].float() / 4.0).permute(0, 2, 3, 1).contiguous()
    got, = db.altcorr_forward(f1, f2, coords, 3)
    want, = ref.altcorr_forward(f1, f2, coords, 3)
    assert relerr(_np(got), _np(want)) < 1e-5
    g = torch.randn_like(want)
    a = db.altcorr_backward(f1, f2, coords, g, 3)
    b = ref.altcorr_backward(f1, f2, coords, g, 3)
    assert relerr(_np(a[0]), _np(b[0])) < 1e-5 and relerr(_np(a[1]), _np(b[1])) < 1e-4
