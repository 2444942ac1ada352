"""CPU checks of the oracle itself (no GPU): its formulas against independent derivations
(finite differences, brute-force bookkeeping, torch fp32 references), so that the checker the
GPU parity tests lean on is itself checked."""
import numpy as np
import pytest
import torch

from helpers import relerr


# ------------------------------------------------------------------ SE3 conventions ----
def test_se3_retraction_matches_matrix_exponential(oracle):
    rng = np.random.default_rng(0)
    xi = rng.normal(0, 0.3, (5, 6))
    t = rng.normal(0, 1, (5, 3))
    q = rng.normal(0, 1, (5, 4))
    q /= np.linalg.norm(q, axis=-1, keepdims=True)
    t1, q1 = oracle.retr_se3(xi, t, q)

    def mat(t_, q_):
        x, y, z, w = q_
        R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                      [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                      [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
        M = np.eye(4)
        M[:3, :3] = R
        M[:3, 3] = t_
        return M

    from scipy.linalg import expm
    for k in range(5):
        tau, phi = xi[k, :3], xi[k, 3:]
        A = np.zeros((4, 4))
        A[:3, :3] = np.array([[0, -phi[2], phi[1]], [phi[2], 0, -phi[0]], [-phi[1], phi[0], 0]])
        A[:3, 3] = tau
        want = expm(A) @ mat(t[k], q[k])       # left-multiplicative retraction, tangent = (tau, phi)
        got = mat(t1[k], q1[k])
        assert np.abs(want - got).max() < 1e-9


def test_relative_pose_and_action(oracle):
    rng = np.random.default_rng(1)
    P = rng.normal(0, 1, (2, 7))
    P[:, 3:] /= np.linalg.norm(P[:, 3:], axis=-1, keepdims=True)
    tij, qij = oracle.rel_se3(P[0, :3], P[0, 3:], P[1, :3], P[1, 3:])
    X = rng.normal(0, 1, 3)
    # a world point seen in camera i then moved by Gij must equal the point seen in camera j
    Xi = oracle._rot(P[0, 3:], X) + P[0, :3]
    Xj = oracle._rot(P[1, 3:], X) + P[1, :3]
    got = oracle.act_se3(tij, qij, np.concatenate([Xi, [1.0]]))[:3]
    assert np.abs(got - Xj).max() < 1e-12


# ---------------------------------------------------------- Jacobians: finite differences ----
def _residuals(oracle, poses, disps, intr, targets, ii, jj):
    """u/v reprojection residuals of every edge/pixel, float64, no MIN_DEPTH gating."""
    fx, fy, cx, cy = intr
    N, ht, wd = disps.shape
    _, _, X0, X1 = oracle._pixel_grid(ht, wd, intr, np.float64)
    tij, qij = oracle._edge_transforms(poses, ii, jj)
    d0 = disps.reshape(N, -1)[ii]
    n = len(ii)
    Xi = np.stack([np.broadcast_to(X0, d0.shape), np.broadcast_to(X1, d0.shape), np.ones_like(d0), d0], -1)
    Xj = oracle.act_se3(tij[:, None], qij[:, None], Xi)
    u = fx * Xj[..., 0] / Xj[..., 2] + cx
    v = fy * Xj[..., 1] / Xj[..., 2] + cy
    tg = targets.reshape(n, 2, -1)
    return tg[:, 0] - u, tg[:, 1] - v


def test_jacobians_match_finite_differences(oracle, synth):
    w = synth.make_window("tiny", seed=3)
    poses = w["poses"].astype(np.float64)
    disps = w["disps"].astype(np.float64)
    intr = w["intrinsics"].astype(np.float64)
    tg = w["targets"].astype(np.float64)
    ii, jj = w["ii"], w["jj"]
    ones = np.ones_like(tg) * 1000.0           # weight 1000 * 0.001 = 1
    Hs, vs, Eii, Eij, Cii, bz = oracle.jacobian_edges(poses, disps, intr, tg, ones, ii, jj, np.float64)
    ru, rv = _residuals(oracle, poses, disps, intr, tg, ii, jj)
    eps = 1e-5
    # gradient wrt pose j of 0.5*sum r^2 is -vs[1]; perturb pose j on the left: T <- Exp(eps*e_k) T
    def cost(p_, e):
        a, b = _residuals(oracle, p_, disps, intr, tg[e:e + 1], ii[e:e + 1], jj[e:e + 1])
        return 0.5 * (a ** 2 + b ** 2).sum()

    for e in (0, 3, len(ii) - 1):
        for k in range(6):
            for which, frame in ((0, ii[e]), (1, jj[e])):
                f = []
                for sgn in (+1.0, -1.0):
                    xi = np.zeros(6)
                    xi[k] = sgn * eps
                    p2 = poses.copy()
                    t, q = oracle.retr_se3(xi[None], p2[frame:frame + 1, :3], p2[frame:frame + 1, 3:])
                    p2[frame, :3], p2[frame, 3:] = t[0], q[0]
                    f.append(cost(p2, e))
                # r(x+dx) ~ r - J dx  =>  sum r J = -(d/dx) 0.5 sum r^2   (central difference)
                num = -(f[0] - f[1]) / (2 * eps)
                assert abs(num - vs[which, e, k]) < 1e-4 * max(1.0, abs(vs[which, e, k])), (e, k, which)
    # depth gradient: bz = sum_uv r * Jz  with Jz = d(proj)/d(disparity)
    ru2, rv2 = _residuals(oracle, poses, disps + eps, intr, tg, ii, jj)
    ru3, rv3 = _residuals(oracle, poses, disps - eps, intr, tg, ii, jj)
    num = -((0.5 * (ru2 ** 2 + rv2 ** 2)) - 0.5 * (ru3 ** 2 + rv3 ** 2)) / (2 * eps)
    assert relerr(num, bz) < 1e-4


def test_hessian_blocks_are_consistent(oracle, synth):
    w = synth.make_window("tiny", seed=4, stereo_edges=2)
    Hs, vs, Eii, Eij, Cii, bz = oracle.jacobian_edges(
        w["poses"], w["disps"], w["intrinsics"], w["targets"], w["weights"], w["ii"], w["jj"], np.float64)
    assert np.allclose(Hs[1], np.transpose(Hs[2], (0, 2, 1)))
    assert np.allclose(Hs[0], np.transpose(Hs[0], (0, 2, 1)))
    stereo = w["ii"] == w["jj"]
    assert stereo.sum() == 2
    # stereo edges contribute to the depth blocks only (src/droid_kernels.cu:367,400)
    assert np.all(Hs[:, stereo] == 0) and np.all(vs[:, stereo] == 0) and np.all(Eij[stereo] == 0)
    assert np.any(Cii[stereo] != 0)


# ------------------------------------------------------------------------ bookkeeping ----
def test_tables_bruteforce(oracle):
    rng = np.random.default_rng(5)
    for trial in range(20):
        N = int(rng.integers(4, 30))
        E = int(rng.integers(0, 60))
        t0 = int(rng.integers(0, N - 1))
        t1 = int(rng.integers(t0 + 1, N + 1))
        ii = rng.integers(0, t1, E)
        jj = rng.integers(0, t1, E)
        tab = oracle.build_tables(ii, jj, t0, t1)
        frames = sorted(set(range(t0, t1)) | set(int(x) for x in ii))
        assert list(tab["kx"]) == frames
        assert all(frames[k] == f for k, f in zip(tab["kk_exp"], tab["ii_exp"]))
        ptr, idx = oracle.csr_by_source(ii, tab["kx"])
        for k, f in enumerate(frames):
            assert list(idx[ptr[k]:ptr[k + 1]]) == [e for e in range(E) if ii[e] == f]
        tri = oracle.schur_triples(tab, t0, t1)
        want = set()
        P = t1 - t0
        rows = [(n, int(tab["jj_exp"][n]) - t0, int(tab["kk_exp"][n])) for n in range(P + E)]
        for n, pa, ka in rows:
            for m, pb, kb in rows:
                if 0 <= pa < P and 0 <= pb < P and ka == kb:
                    want.add((pa, pb, n, m, ka))
        assert set(map(tuple, tri.tolist())) == want and len(tri) == len(want)


# ------------------------------------------------------- BA: dense normal equations ----
def test_schur_path_equals_full_normal_equations(oracle, synth):
    """The reduced system + back-substitution must solve the full (pose, depth) normal equations
    whenever the reference's quirk is inactive (no row of E belongs to pose index 0)."""
    w = synth.make_window("tiny", seed=6)
    t0, t1 = 1, w["t1"]
    keep = (w["jj"] != t0) & (w["ii"] != t0)       # isolate the first window pose
    ii, jj = w["ii"][keep], w["jj"][keep]
    tg, wt = w["targets"][keep], w["weights"][keep]
    tab = oracle.build_tables(ii, jj, t0, t1)
    eta = np.full((len(tab["kx"]), w["ht"], w["wd"]), 1e-3)
    p2, d2 = w["poses"].astype(np.float64), w["disps"].astype(np.float64)
    tr = []
    ep = 0.1
    oracle.ba(p2, d2, w["intrinsics"], np.zeros_like(d2), tg, wt, eta, ii, jj, t0, t1, 1, 0.0, ep,
              False, dtype=np.float64, trace=tr)
    T = tr[0]
    E, Q, ww = T["E"], T["Q"], T["w"]
    P, HW, K = t1 - t0, w["ht"] * w["wd"], len(tab["kx"])
    A, b = oracle.assemble_pose_system(T["Hs"], T["vs"], ii, jj, t0, t1)
    # full system: [A  Ebig; Ebig^T  C] [dx; dz] = [b; w]
    Ebig = np.zeros((6 * P, K * HW))
    for n in range(E.shape[0]):
        p = tab["jj_exp"][n] - t0
        if p >= 0:
            k = tab["kk_exp"][n]
            Ebig[6 * p:6 * p + 6, k * HW:(k + 1) * HW] += E[n]
    Cd = (1.0 / Q).reshape(-1)
    S = A - (Ebig / Cd) @ Ebig.T
    assert relerr(S, T["H"]) < 1e-9
    assert relerr(b - Ebig @ (ww.reshape(-1) / Cd), T["v"]) < 1e-9
    full = np.block([[A + float(np.float32(ep)) * np.eye(6 * P), Ebig], [Ebig.T, np.diag(Cd)]])
    sol = np.linalg.solve(full, np.concatenate([b, ww.reshape(-1)]))
    assert all(np.all(E[n] == 0) for n in range(E.shape[0]) if tab["jj_exp"][n] - t0 == 0)
    assert relerr(T["dx"], sol[:6 * P]) < 1e-5
    assert relerr(T["dz"].reshape(-1), sol[6 * P:]) < 1e-5


def test_ba_converges_and_motion_only(oracle, synth):
    w = synth.make_window("c1", seed=0)
    poses, disps = w["poses"].copy(), w["disps"].copy()
    err0 = np.abs(poses[:, :3] - w["poses_gt"][:, :3]).max()
    oracle.ba(poses, disps, w["intrinsics"], w["disps_sens"], w["targets"], w["weights"], w["eta"],
              w["ii"], w["jj"], w["t0"], w["t1"], 3, 1e-4, 0.1, False, dtype=np.float32)
    assert np.abs(poses[:, :3] - w["poses_gt"][:, :3]).max() < 0.4 * err0
    assert np.array_equal(poses[0], w["poses"][0])            # frame 0 is outside [t0,t1)
    poses2, disps2 = w["poses"].copy(), w["disps"].copy()
    dx, dz = oracle.ba(poses2, disps2, w["intrinsics"], w["disps_sens"], w["targets"], w["weights"], w["eta"],
                       w["ii"], w["jj"], w["t0"], w["t1"], 2, 1e-4, 0.1, True, dtype=np.float32)
    assert dz is None and np.array_equal(disps2, w["disps"]) and not np.array_equal(poses2, w["poses"])


def test_failed_factorisation_gives_zero_update(oracle):
    A = np.array([[1.0, 2.0], [2.0, 1.0]])      # indefinite
    dx, ok = oracle.damped_solve(A, np.ones(2), 0.0, 0.0)
    assert not ok and np.all(dx == 0)


def test_bacore_equals_ba_with_alpha(oracle, synth):
    """BACore.hessian + an external solve + retract is `ba` with alpha = 0.001."""
    w = synth.make_window("tiny", seed=7, sensor_frac=0.3)
    args = (w["intrinsics"], w["disps_sens"], w["targets"], w["weights"], w["eta"], w["ii"], w["jj"],
            w["t0"], w["t1"])
    p1, d1 = w["poses"].copy(), w["disps"].copy()
    oracle.ba(p1, d1, *args, 1, 1e-4, 0.1, False, dtype=np.float32, alpha=0.001)
    p2, d2 = w["poses"].copy(), w["disps"].copy()
    core = oracle.BACore()
    core.init(p2, d2, *args, 2, 1e-4, 0.1, False)
    n = 6 * (w["t1"] - w["t0"])
    H, v = np.zeros((n, n)), np.zeros(n)
    core.hessian(H, v)
    dx, _ = oracle.damped_solve(H, v, 1e-4, 0.1)
    core.retract(dx.astype(np.float64))
    assert np.allclose(p1, p2, atol=1e-6) and np.allclose(d1, d2, atol=1e-6)


# ------------------------------------------------------------------------ correlation ----
def test_volume_pyramid_vs_torch(oracle, synth):
    f = synth.make_features("tiny", seed=1, n_frames=3)
    f1, f2 = f[[0, 1]], f[[1, 2]]
    pyr = oracle.corr_pyramid(f1, f2, 4)
    a = torch.from_numpy(f1).float().reshape(2, 128, -1) / 4.0
    b = torch.from_numpy(f2).float().reshape(2, 128, -1) / 4.0
    ht, wd = f.shape[2:]
    corr = torch.matmul(a.transpose(1, 2), b).half().reshape(2 * ht * wd, 1, ht, wd)
    for l in range(4):
        want = corr.view(2, ht, wd, ht >> l, wd >> l).numpy()
        assert pyr[l].shape == want.shape
        assert np.abs(pyr[l].astype(np.float32) - want.astype(np.float32)).max() <= 2e-3 * max(1, np.abs(want).max())
        corr = torch.nn.functional.avg_pool2d(corr.float(), 2, stride=2).half()


def test_pool_floor_semantics(oracle):
    v = np.arange(2 * 5 * 7, dtype=np.float16).reshape(2, 5, 7)
    p = oracle.avg_pool2(v)
    assert p.shape == (2, 2, 3)
    want = torch.nn.functional.avg_pool2d(torch.from_numpy(v).float()[:, None], 2, stride=2)[:, 0].half().numpy()
    assert np.array_equal(p, want)


def _bilinear_ref(vol, coords, r):
    n, h1, w1, h2, w2 = vol.shape
    rd = 2 * r + 1
    out = np.zeros((n, rd, rd, h1, w1))
    for e in range(n):
        for y in range(h1):
            for x in range(w1):
                x0, y0 = coords[e, 0, y, x], coords[e, 1, y, x]
                for a in range(rd):
                    for b in range(rd):
                        xs, ys = x0 - r + a, y0 - r + b
                        fx, fy = int(np.floor(xs)), int(np.floor(ys))
                        dx, dy = xs - fx, ys - fy
                        acc = 0.0
                        for (yy, xx, ww) in ((fy, fx, (1 - dx) * (1 - dy)), (fy, fx + 1, dx * (1 - dy)),
                                             (fy + 1, fx, (1 - dx) * dy), (fy + 1, fx + 1, dx * dy)):
                            if 0 <= yy < h2 and 0 <= xx < w2:
                                acc += ww * vol[e, y, x, yy, xx]
                        out[e, a, b, y, x] = acc
    return out


def test_lookup_is_zero_padded_bilinear_sampling(oracle):
    rng = np.random.default_rng(8)
    vol = rng.normal(0, 1, (2, 3, 4, 6, 5)).astype(np.float32)
    coords = rng.uniform(-3, 8, (2, 2, 3, 4)).astype(np.float32)
    got = oracle.corr_index_forward(vol, coords, 3)
    want = _bilinear_ref(vol.astype(np.float64), coords.astype(np.float64), 3)
    assert np.abs(got - want).max() < 1e-5
    got16 = oracle.corr_index_forward(vol.astype(np.float16), coords, 3)
    assert np.abs(got16.astype(np.float64) - want).max() < 6 * 2.0 ** -11 * np.abs(vol).max() * 1.5


def test_lookup_backward_is_adjoint(oracle):
    rng = np.random.default_rng(9)
    shape = (2, 3, 4, 6, 5)
    vol = rng.normal(0, 1, shape).astype(np.float32)
    coords = rng.uniform(-3, 8, (2, 2, 3, 4)).astype(np.float32)
    g = rng.normal(0, 1, (2, 7, 7, 3, 4)).astype(np.float32)
    fwd = oracle.corr_index_forward(vol, coords, 3)
    bwd = oracle.corr_index_backward(shape, coords, g, 3)
    assert abs((fwd.astype(np.float64) * g).sum() - (bwd.astype(np.float64) * vol).sum()) < 1e-3


def test_altcorr_equals_volume_lookup(oracle, synth):
    """C4: the on-the-fly path must give the values of volume + lookup (fp32)."""
    rng = np.random.default_rng(10)
    B, H, W, C = 2, 6, 8, 32
    f1 = rng.normal(0, 1, (B, H, W, C)).astype(np.float32)
    f2 = rng.normal(0, 1, (B, H, W, C)).astype(np.float32)
    coords = rng.uniform(-2, 9, (B, 1, H, W, 2)).astype(np.float32)
    alt = oracle.altcorr_forward(f1, f2, coords, 3)
    vol = np.einsum("bhwc,byxc->bhwyx", f1.astype(np.float64), f2.astype(np.float64)).astype(np.float32)
    look = oracle.corr_index_forward(vol, np.ascontiguousarray(coords[:, 0].transpose(0, 3, 1, 2)), 3)
    # lookup output is [n, x-off, y-off, h, w]; altcorr channel = y-off + 7*x-off: same flattening
    assert relerr(alt[:, 0], look.reshape(B, 49, H, W)) < 1e-5
    g = rng.normal(0, 1, alt.shape).astype(np.float32)
    g1, g2, gc = oracle.altcorr_backward(f1, f2, coords, g, 3)
    assert np.all(gc == 0)
    lhs = (alt.astype(np.float64) * g).sum()
    assert abs(lhs - (g1.astype(np.float64) * f1).sum()) < 1e-2 * max(1, abs(lhs))
    assert abs(lhs - (g2.astype(np.float64) * f2).sum()) < 1e-2 * max(1, abs(lhs))


# --------------------------------------------------------------------------- geometry ----
def test_projmap_frame_distance_iproj(oracle, synth):
    w = synth.make_window("tiny", seed=11, near_frac=0.1)
    c, valid = oracle.projmap(w["poses"], w["disps"], w["intrinsics"], w["ii"], w["jj"])
    assert c.shape == (len(w["ii"]), w["ht"], w["wd"], 3) and np.all(c[..., 2] == 0)
    assert 0 < valid.mean() < 1
    d = oracle.frame_distance(w["poses"], w["disps"], w["intrinsics"], w["ii"], w["jj"], 0.3)
    assert d.shape == (len(w["ii"]),) and np.all(d > 0)
    # identical frames with zero motion: distance 0
    p = w["poses"].copy()
    p[1] = p[0]
    d0 = oracle.frame_distance(p, np.full_like(w["disps"], 0.5), w["intrinsics"], np.array([0]), np.array([1]), 0.3)
    assert abs(d0[0]) < 1e-4
    pts = oracle.iproj(w["poses"], w["disps"], w["intrinsics"])
    assert pts.shape == w["disps"].shape + (3,)
    cnt = oracle.depth_filter(w["poses_gt"], w["disps_gt"], w["intrinsics"], np.array([1, 3]),
                              np.array([0.05, 0.05], np.float32))
    assert cnt.shape == (2, w["ht"], w["wd"]) and cnt.max() <= 6 and cnt.mean() > 0.5


def test_ba2gtsam_blockwise(oracle):
    """oracle.ba2gtsam (BA2GTSAM, dbaf/depth_video.py:20-29) against the block-wise form A^T H_IK A, A^T v_I that the
    device kernel uses, and its invariants: symmetry is preserved, A = -I leaves H and flips v."""
    rng = np.random.default_rng(3)
    P = 5
    G = rng.normal(0, 1, (6 * P, 6 * P))
    H, v, A = G @ G.T, rng.normal(0, 1, 6 * P), rng.normal(0, 1, (6, 6))
    Hg, vg = oracle.ba2gtsam(H, v, A)
    for I in range(P):
        assert np.allclose(vg[6 * I:6 * I + 6], A.T @ v[6 * I:6 * I + 6])
        for K in range(P):
            assert np.allclose(Hg[6 * I:6 * I + 6, 6 * K:6 * K + 6], A.T @ H[6 * I:6 * I + 6, 6 * K:6 * K + 6] @ A)
    assert np.allclose(Hg, Hg.T)
    H2, v2 = oracle.ba2gtsam(H, v, -np.eye(6))
    assert np.allclose(H2, H) and np.allclose(v2, -v)


def test_ba2gtsam_matrix_is_the_row_swapped_negative_adjoint(db):
    """A = rows-swapped -Adjoint(Tbc^-1) (dbaf/depth_video.py:21-23).  The Adjoint is pinned by its defining property
    T hat(xi) T^-1 = hat(Adjoint(T) xi) for GTSAM's (omega, v) tangent order, checked on random poses."""
    rng = np.random.default_rng(0)

    def hat(xi):
        w, v = xi[:3], xi[3:]
        M = np.zeros((4, 4))
        M[:3, :3] = [[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]]
        M[:3, 3] = v
        return M

    for _ in range(5):
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        x, y, z, w = q
        R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                      [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                      [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
        T = np.eye(4)
        T[:3, :3], T[:3, 3] = R, rng.normal(size=3)
        A = db.ba2gtsam_matrix(T)
        A0 = -np.concatenate([A[3:6], A[0:3]], axis=0)            # undo the swap and the sign: Adjoint(T^-1)
        Ti = np.linalg.inv(T)
        for _ in range(3):
            xi = rng.normal(size=6)
            assert np.allclose(Ti @ hat(xi) @ T, hat(A0 @ xi), atol=1e-12)


@pytest.mark.parametrize("m,levels", [(1, 1), (2, 2), (3, 2), (5, 3), (7, 3), (9, 4), (12, 4), (15, 4)])
def test_block_cyclic_reduction_matches_dense_solve(oracle, m, levels):
    """The elimination order of the domain-decomposed banded solve's separator system (csrc/ba_spike.cuh) restated
    in numpy: same solution as a dense solve, log2 depth."""
    rng = np.random.default_rng(m)
    ns = 6
    n = m * ns
    M = rng.normal(size=(n, n))
    A = M @ M.T
    for i in range(m):
        for j in range(m):
            if abs(i - j) > 1:
                A[i * ns:(i + 1) * ns, j * ns:(j + 1) * ns] = 0.0
    A += 2 * n * np.eye(n)
    b = rng.normal(size=n)
    D = {s: A[(s - 1) * ns:s * ns, (s - 1) * ns:s * ns] for s in range(1, m + 1)}
    C = {a: A[a * ns:(a + 1) * ns, (a - 1) * ns:a * ns] for a in range(1, m)}
    y = {s: b[(s - 1) * ns:s * ns] for s in range(1, m + 1)}
    x, lev = oracle.block_tridiagonal_cyclic_reduction(D, C, y)
    got = np.concatenate([x[s] for s in range(1, m + 1)])
    want = np.linalg.solve(A, b)
    assert np.abs(got - want).max() <= 1e-12 * np.abs(want).max()
    assert lev == levels


def test_banded_solve_plan_host_side(db):
    """Host-side plan of dbaf_ba_solve_banded (no device work): launch counts per path, strip size of the band."""
    assert db.banded_solve_launches(11, 0) == 1                     # dense shared-memory solve
    assert db.banded_solve_launches(40, 20) == 1                    # one sweep: chain too short to cut
    assert db.banded_solve_launches(100, 20) == 3                   # two sweeps, separator, back-substitutions
    # 16 domains, 15 separators: sweeps, spike + Gram, 3 x (eliminate, update), last block, 3 x back, apply, back
    assert db.banded_solve_launches(999, 20) == 1 + 2 + 6 + 1 + 3 + 1 + 1
    from droid_backends import _lib
    assert _lib.lib.dbaf_ba_band_strip_doubles(999, 20) == 5994 * 6 * 41 + 5994
    assert _lib.lib.dbaf_ba_layout_bytes() == _lib.ctypes.sizeof(_lib.BaLayout)
