"""Golden vectors produced by the UNMODIFIED reference CUDA extension on a B200
(tests/golden/make_golden.py -> tests/golden/ref_*.npz).

CPU half (`-m "not gpu"`): pins the numpy oracle against the reference's own outputs.
GPU half: the CUDA path against the same vectors.
"""
import hashlib
import importlib.util
import os

import numpy as np
import pytest

from helpers import ba_args, relerr, to_torch

HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
TOL = 1e-4


def _cases():
    # parse GOLDEN_CASES without importing torch-heavy module state twice
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    return mg


MG = _cases()


def _load(name):
    path = os.path.join(HERE, "golden", f"ref_{name}.npz")
    if not os.path.exists(path):
        pytest.skip(f"{path} missing")
    return np.load(path)


def _window(synth, name):
    win = synth.make_window(**MG.GOLDEN_CASES[name])
    return win


@pytest.mark.parametrize("name", ["tiny", "c1"])
def test_inputs_are_reproducible(synth, name):
    g = _load(name)
    win = _window(synth, name)
    assert bytes(g["checksum"]).hex() == MG.checksum(win), "synthetic generator drifted from the golden inputs"


def _slack(ref, truth):
    # pose/depth updates solve an ill-conditioned system: allow what the reference itself deviates
    # from the float64 solution (x3), never less than 1e-4
    return max(TOL, 3 * relerr(ref, truth))


@pytest.mark.parametrize("name", ["tiny", "c1"])
def test_oracle_ba_matches_reference(oracle, synth, name):
    g = _load(name)
    w = _window(synth, name)
    common = (w["intrinsics"], w["disps_sens"], w["targets"], w["weights"], w["eta"], w["ii"], w["jj"], w["t0"], w["t1"])
    n = 6 * (w["t1"] - w["t0"])
    # reduced camera system exported by BACore.hessian: float32 kernels, float64 assembly
    for dt in (np.float32, np.float64):
        core = oracle.BACore()
        p, d = w["poses"].astype(dt), w["disps"].astype(dt)
        core.init(p, d, *common, 2, 1e-4, 0.1, False, dtype=dt)
        H, v = np.zeros((n, n)), np.zeros(n)
        core.hessian(H, v)
        assert relerr(H, g["core_H"]) < TOL and relerr(v, g["core_v"]) < TOL
        dx, dz = core.retract(g["core_dx_in"])
        assert relerr(dx, g["core_dx"]) < 1e-6 and relerr(dz, g["core_dz"]) < TOL
        assert relerr(p, g["core_poses"]) < 1e-5 and relerr(d, g["core_disps"]) < TOL
    # one and two full iterations, and motion-only
    for iters, key in ((1, "ba1"), (2, "ba")):
        p64, d64 = w["poses"].astype(np.float64), w["disps"].astype(np.float64)
        dx64, dz64 = oracle.ba(p64, d64, *common, iters, 1e-4, 0.1, False, dtype=np.float64)
        p32, d32 = w["poses"].copy(), w["disps"].copy()
        dx32, dz32 = oracle.ba(p32, d32, *common, iters, 1e-4, 0.1, False, dtype=np.float32)
        # the reference must agree with the float64 restatement to within float32 noise, measured
        # by the restatement's own float32 mode (the solve amplifies rounding by cond(H) ~ 1e4..1e6)
        assert relerr(g[key + "_dx"], dx64) < _slack(dx32, dx64)
        assert relerr(g[key + "_dz"], dz64) < _slack(dz32, dz64)
        assert relerr(g[key + "_poses"], p64) < TOL
        assert relerr(g[key + "_disps"], d64) < _slack(d32, d64)
    p64, d64 = w["poses"].astype(np.float64), w["disps"].astype(np.float64)
    dx64, _ = oracle.ba(p64, d64, *common, 2, 1e-4, 0.1, True, dtype=np.float64)
    p32, d32 = w["poses"].copy(), w["disps"].copy()
    dx32, _ = oracle.ba(p32, d32, *common, 2, 1e-4, 0.1, True, dtype=np.float32)
    assert relerr(g["mo_dx"], dx64) < _slack(dx32, dx64) and relerr(g["mo_poses"], p64) < TOL


@pytest.mark.parametrize("name", ["tiny", "c1"])
def test_oracle_geometry_matches_reference(oracle, synth, name):
    g = _load(name)
    w = _window(synth, name)
    c, va = oracle.projmap(w["poses"], w["disps"], w["intrinsics"], w["ii"][:4], w["jj"][:4], np.float64)
    assert (va != g["projmap_valid"]).mean() < 1e-3
    ok = g["projmap_valid"][..., 0] > 0
    assert relerr(c[ok], g["projmap_coords"][ok]) < 1e-5
    fd = oracle.frame_distance(w["poses"], w["disps"], w["intrinsics"], w["ii"], w["jj"], 0.3, np.float64)
    far = g["frame_distance"] > 999
    assert np.array_equal(fd > 999, far) and relerr(fd[~far], g["frame_distance"][~far]) < 1e-4
    assert relerr(oracle.iproj(w["poses"], w["disps"], w["intrinsics"], np.float64)[:3], g["iproj"]) < 1e-5
    m = len(g["depth_filter"])
    cnt = oracle.depth_filter(w["poses"], w["disps"], w["intrinsics"], np.arange(m), np.full(m, 0.05, np.float32))
    assert (cnt != g["depth_filter"]).mean() < 2e-3


def test_oracle_correlation_matches_reference(oracle, synth):
    g = _load("tiny")
    case = MG.GOLDEN_CASES["tiny"]
    w = _window(synth, "tiny")
    f1, f2, coords = MG.corr_inputs(case, w)
    # volume: fp32-accumulated product, one rounding (see test_gpu_parity for the tolerance)
    vol0 = oracle.corr_volume(f1, f2)
    d = np.abs(vol0[:, ::3, ::3].astype(np.float32) - g["vol0_sample"].astype(np.float32))
    assert np.all(d <= 2.0 ** -10 * np.abs(g["vol0_sample"].astype(np.float32)) + 8e-6)
    # pooling: bit-exact given the reference's own previous level
    assert np.array_equal(oracle.avg_pool2(g["vol1"]).view(np.uint16), g["vol2"].view(np.uint16))
    assert np.array_equal(oracle.avg_pool2(g["vol2"]).view(np.uint16), g["vol3"].view(np.uint16))
    # lookup: bit-exact (fp16 rounding sequence) on the reference's own volumes
    for l in (1, 2, 3):
        got = oracle.corr_index_forward(g[f"vol{l}"], coords / np.float32(2 ** l), 3)
        assert np.array_equal(got.view(np.uint16), g[f"lookup{l}"].view(np.uint16)), l
    vol2f = g["vol2"].astype(np.float32)
    c2 = coords / np.float32(4)
    assert relerr(oracle.corr_index_forward(vol2f, c2, 3), g["lookup2_f32"]) < 1e-6
    gr = np.random.default_rng(3).normal(0, 1, g["lookup2_f32"].shape).astype(np.float32)
    assert relerr(oracle.corr_index_backward(vol2f.shape, c2, gr, 3), g["lookup2_bwd"]) < 1e-6
    # altcorr
    af1 = np.ascontiguousarray((f1.astype(np.float32) / 4.0).transpose(0, 2, 3, 1))
    af2 = np.ascontiguousarray((f2.astype(np.float32) / 4.0).transpose(0, 2, 3, 1))
    ac = np.ascontiguousarray(coords.transpose(0, 2, 3, 1)[:, None])
    assert relerr(oracle.altcorr_forward(af1, af2, ac, 3, np.float64), g["altcorr"]) < 1e-5
    ga = np.random.default_rng(4).normal(0, 1, g["altcorr"].shape).astype(np.float32)
    g1, g2, _ = oracle.altcorr_backward(af1, af2, ac, ga, 3, np.float64)
    assert relerr(g1[:1], g["altcorr_g1"]) < 1e-5 and relerr(g2[:1], g["altcorr_g2"]) < 1e-4


# ------------------------------------------------------------------------------- GPU half
@pytest.mark.gpu
@pytest.mark.parametrize("name", ["tiny", "c1"])
def test_cuda_ba_matches_golden(dev, db, oracle, synth, name):
    import torch
    g = _load(name)
    w = _window(synth, name)
    n = 6 * (w["t1"] - w["t0"])
    t = to_torch(w, dev)
    core = db.BACore()
    core.init(*ba_args(t), w["t0"], w["t1"], 2, 1e-4, 0.1, False)
    H = torch.zeros(n, n, dtype=torch.float64)
    v = torch.zeros(n, dtype=torch.float64)
    core.hessian(H, v)
    assert relerr(H.numpy(), g["core_H"]) < TOL and relerr(v.numpy(), g["core_v"]) < TOL
    dx, dz = core.retract(torch.from_numpy(g["core_dx_in"]))
    assert relerr(dx.cpu().numpy(), g["core_dx"]) < 1e-6 and relerr(dz.cpu().numpy(), g["core_dz"]) < TOL
    assert relerr(t["poses"].cpu().numpy(), g["core_poses"]) < 1e-6
    assert relerr(t["disps"].cpu().numpy(), g["core_disps"]) < TOL
    common = (w["intrinsics"], w["disps_sens"], w["targets"], w["weights"], w["eta"], w["ii"], w["jj"], w["t0"], w["t1"])
    for iters, key in ((1, "ba1"), (2, "ba")):
        t = to_torch(w, dev)
        dx, dz = db.ba(*ba_args(t), w["t0"], w["t1"], iters, 1e-4, 0.1, False)
        p64, d64 = w["poses"].astype(np.float64), w["disps"].astype(np.float64)
        dx64, dz64 = oracle.ba(p64, d64, *common, iters, 1e-4, 0.1, False, dtype=np.float64)
        assert relerr(dx.cpu().numpy(), g[key + "_dx"]) < _slack(g[key + "_dx"], dx64)
        assert relerr(dz.cpu().numpy(), g[key + "_dz"]) < _slack(g[key + "_dz"], dz64)
        assert relerr(t["poses"].cpu().numpy(), g[key + "_poses"]) < _slack(g[key + "_poses"], p64)
        assert relerr(t["disps"].cpu().numpy(), g[key + "_disps"]) < _slack(g[key + "_disps"], d64)
    t = to_torch(w, dev)
    dx, dz = db.ba(*ba_args(t), w["t0"], w["t1"], 2, 1e-4, 0.1, True)
    assert dz is None and relerr(t["poses"].cpu().numpy(), g["mo_poses"]) < TOL


@pytest.mark.gpu
def test_cuda_geometry_and_correlation_match_golden(dev, db, oracle, synth):
    import torch
    g = _load("tiny")
    case = MG.GOLDEN_CASES["tiny"]
    w = _window(synth, "tiny")
    t = to_torch(w, dev)
    c, va = db.projmap(t["poses"], t["disps"], t["intrinsics"], t["ii"][:4].contiguous(), t["jj"][:4].contiguous())
    assert (va.cpu().numpy() != g["projmap_valid"]).mean() < 1e-3
    assert relerr(c.cpu().numpy(), g["projmap_coords"]) < 1e-5
    fd = db.frame_distance(t["poses"], t["disps"], t["intrinsics"], t["ii"], t["jj"], 0.3).cpu().numpy()
    assert relerr(fd, g["frame_distance"]) < 1e-4
    assert relerr(db.iproj(t["poses"], t["disps"], t["intrinsics"]).cpu().numpy()[:3], g["iproj"]) < 1e-6
    m = len(g["depth_filter"])
    cnt = db.depth_filter(t["poses"], t["disps"], t["intrinsics"], torch.arange(m, device=dev),
                          torch.full((m,), 0.05, device=dev))
    assert (cnt.cpu().numpy() != g["depth_filter"]).mean() < 1e-3
    f1, f2, coords = MG.corr_inputs(case, w)
    ct = torch.from_numpy(coords).to(dev)
    pyr = db.corr_volume_pyramid(torch.from_numpy(f1).to(dev), torch.from_numpy(f2).to(dev), 4)
    d = (pyr[0][:, ::3, ::3].float().cpu().numpy() - g["vol0_sample"].astype(np.float32))
    assert np.all(np.abs(d) <= 2.0 ** -10 * np.abs(g["vol0_sample"].astype(np.float32)) + 8e-6)
    scale = float(np.abs(g["vol0_sample"].astype(np.float32)).max())
    for l in (1, 2, 3):
        want = g[f"vol{l}"].astype(np.float32)
        mine = pyr[l].cpu().numpy()
        # a pooled value inherits up to one fp16 ulp of each of its (larger) inputs: bound it by the
        # ulp of the largest level-0 value; the pooling arithmetic itself is bit-exact
        assert np.all(np.abs(mine.astype(np.float32) - want) <= (l + 1) * 2.0 ** -10 * scale)
        assert np.array_equal(oracle.avg_pool2(pyr[l - 1].cpu().numpy()).view(np.uint16), mine.view(np.uint16))
        got, = db.corr_index_forward(torch.from_numpy(g[f"vol{l}"]).to(dev), (ct / 2 ** l).contiguous(), 3)
        assert np.array_equal(got.cpu().numpy().view(np.uint16), g[f"lookup{l}"].view(np.uint16)), l
    af1 = (torch.from_numpy(f1).to(dev).float() / 4.0).permute(0, 2, 3, 1).contiguous()
    af2 = (torch.from_numpy(f2).to(dev).float() / 4.0).permute(0, 2, 3, 1).contiguous()
    ac = ct.permute(0, 2, 3, 1)[:, None].contiguous()
    alt, = db.altcorr_forward(af1, af2, ac, 3)
    assert relerr(alt.cpu().numpy(), g["altcorr"]) < 1e-5


# ----------------------------------------------------------------- reproject
def _reproject_cases():
    g = np.load(os.path.join(HERE, "golden", "ref_reproject.npz"))
    for name in ("tiny", "c3"):
        yield name, {k[len(name) + 1:]: g[k] for k in g.files if k.startswith(name + "_")}


def test_oracle_reproject_matches_reference_python(oracle):
    """oracle.reproject against dbaf/geom/projective_ops.py run on CPU (tests/golden/make_golden_reproject.py):
    coordinates to 2e-3 px where the point is in front of the camera (fp32 evaluation order differs), validity
    exact away from the 0.2 threshold; stereo edges and per-frame intrinsics are in the fixture."""
    for name, c in _reproject_cases():
        assert (c["ii"] == c["jj"]).any()
        for dt, tol in ((np.float64, 2e-3), (np.float32, 4e-3)):
            coords, valid = oracle.reproject(c["poses"], c["disps"], c["intr"], c["ii"], c["jj"], dtype=dt)
            front = c["valid"][..., 0] > 0
            assert np.abs(coords - c["coords"])[front].max() < tol, (name, dt)
            assert (valid != c["valid"]).mean() < 1e-3, name


@pytest.mark.gpu
def test_reproject_matches_golden_and_oracle(dev, db, oracle):
    import torch
    for name, c in _reproject_cases():
        t = {k: torch.from_numpy(v).to(dev) for k, v in c.items()}
        coords, valid = db.reproject(t["poses"], t["disps"], t["intr"], t["ii"], t["jj"])
        assert coords.shape == (1,) + c["coords"].shape and valid.shape == (1,) + c["valid"].shape
        front = c["valid"][..., 0] > 0
        got = coords[0].cpu().numpy()
        assert np.abs(got - c["coords"])[front].max() < 4e-3, name
        assert (valid[0].cpu().numpy() != c["valid"]).mean() < 1e-3
        want, wv = oracle.reproject(c["poses"], c["disps"], c["intr"], c["ii"], c["jj"], dtype=np.float64)
        assert np.abs(got - want)[front].max() < 4e-3
        # shared intrinsics: a [4] vector gives the same as the same row repeated
        K1 = t["intr"][0].contiguous()
        a, _ = db.reproject(t["poses"], t["disps"], K1, t["ii"], t["jj"])
        b, _ = db.reproject(t["poses"], t["disps"], K1[None].repeat(t["poses"].shape[0], 1).contiguous(), t["ii"], t["jj"])
        assert torch.equal(a, b)
