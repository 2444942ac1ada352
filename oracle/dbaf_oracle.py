"""CPU restatement (numpy) of DBA-Fusion's `droid_backends` operator family.

TEST INFRASTRUCTURE ONLY.  Nothing under ``dba-fusion_b200/`` may import this
file; only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` use it, and only as the checker.

Every function cites the reference lines it restates (paths relative to the
reference checkout, i.e. ``src/droid_kernels.cu`` means
``/root/reference/src/droid_kernels.cu``).

Pinning status: DBA-Fusion ships no tests or golden vectors for this path
(SURVEY.md section 4), so this restatement is pinned against outputs of the
reference CUDA extension itself, rebuilt for sm_100a by ``oracle/build_ref.py``
and run on the B200 box (``tests/test_ref_parity.py``; fixtures produced by
``tests/golden/make_golden.py``).  The fixtures are committed
(``tests/golden/ref_tiny.npz``, ``ref_c1.npz``) and ``tests/test_golden.py``
checks this file against them on CPU: parity is pinned.

All arithmetic is done in ``dtype`` (float64 = "truth", float32 = same
precision as the reference's kernels; numpy never fuses multiply-adds, the GPU
does, so float32 results agree with the device to rounding, not bit for bit).
Integer bookkeeping is exact.
"""
from __future__ import annotations

import numpy as np

MIN_DEPTH = 0.25            # src/droid_kernels.cu:29
STEREO_BASELINE = -0.1      # src/droid_kernels.cu:265


# --------------------------------------------------------------------------
# SE3 helpers (quaternion xyzw, pose = world->camera, tangent = (tau, phi))
# --------------------------------------------------------------------------
def _rot(q, X):
    """Rotate X[...,3] by unit quaternion q[...,4] (x,y,z,w).  src/droid_kernels.cu:61-71."""
    qv = q[..., :3]
    uv = 2.0 * np.cross(qv, X)
    return X + q[..., 3:4] * uv + np.cross(qv, uv)


def rel_se3(ti, qi, tj, qj):
    """Gij = Tj * Ti^-1.  src/droid_kernels.cu:99-110."""
    xi, yi, zi, wi = (qi[..., k] for k in range(4))
    xj, yj, zj, wj = (qj[..., k] for k in range(4))
    qij = np.stack([
        -wj * xi + xj * wi - yj * zi + zj * yi,
        -wj * yi + yj * wi - zj * xi + xj * zi,
        -wj * zi + zj * wi - xj * yi + yj * xi,
        wj * wi + xj * xi + yj * yi + zj * zi], axis=-1)
    tij = tj - _rot(qij, ti)
    return tij, qij


def act_se3(t, q, X4):
    """Y = (R X[:3] + X[3] t, X[3]).  src/droid_kernels.cu:73-80."""
    Y3 = _rot(q, X4[..., :3]) + X4[..., 3:4] * t
    return np.concatenate([Y3, X4[..., 3:4]], axis=-1)


def adj_se3(t, q, X6):
    """Dual-adjoint used for Ji.  src/droid_kernels.cu:82-97."""
    qinv = q * np.array([-1, -1, -1, 1], dtype=q.dtype)
    a = _rot(qinv, X6[..., :3])
    b = _rot(qinv, X6[..., 3:])
    u = np.cross(X6[..., :3], t)
    return np.concatenate([a, b + _rot(qinv, u)], axis=-1)


def exp_so3(phi):
    """src/droid_kernels.cu:113-135 (small-angle branch theta^2 < 1e-8)."""
    th2 = np.sum(phi * phi, axis=-1)
    th4 = th2 * th2
    th = np.sqrt(th2)
    small = th2 < 1e-8
    ths = np.where(small, 1.0, th)
    imag = np.where(small, 0.5 - th2 / 48.0 + th4 / 3840.0, np.sin(0.5 * ths) / ths)
    real = np.where(small, 1.0 - th2 / 8.0 + th4 / 384.0, np.cos(0.5 * ths))
    return np.concatenate([imag[..., None] * phi, real[..., None]], axis=-1).astype(phi.dtype)


def exp_se3(xi):
    """src/droid_kernels.cu:150-178 (V-matrix terms only when theta > 1e-4)."""
    tau, phi = xi[..., :3], xi[..., 3:]
    q = exp_so3(phi)
    th2 = np.sum(phi * phi, axis=-1)
    th = np.sqrt(th2)
    big = th > 1e-4
    th2s = np.where(big, th2, 1.0)
    ths = np.where(big, th, 1.0)
    a = np.where(big, (1 - np.cos(ths)) / th2s, 0.0)[..., None]
    b = np.where(big, (ths - np.sin(ths)) / (ths * th2s), 0.0)[..., None]
    c1 = np.cross(phi, tau)
    c2 = np.cross(phi, c1)
    t = tau + a * c1 + b * c2
    return t.astype(xi.dtype), q


def retr_se3(xi, t, q):
    """T <- Exp(xi) * T.  src/droid_kernels.cu:922-940."""
    dt, dq = exp_se3(xi)
    x, y, z, w = (q[..., k] for k in range(4))
    dx, dy, dz, dw = (dq[..., k] for k in range(4))
    q1 = np.stack([
        dw * x + dx * w + dy * z - dz * y,
        dw * y + dy * w + dz * x - dx * z,
        dw * z + dz * w + dx * y - dy * x,
        dw * w - dx * x - dy * y - dz * z], axis=-1)
    t1 = _rot(dq, t) + dt
    return t1, q1


def _edge_transforms(poses, ii, jj, stereo_special=True):
    ti, qi = poses[ii, :3], poses[ii, 3:]
    tj, qj = poses[jj, :3], poses[jj, 3:]
    tij, qij = rel_se3(ti, qi, tj, qj)
    if stereo_special:
        st = ii == jj
        tij[st] = np.array([STEREO_BASELINE, 0, 0], dtype=poses.dtype)
        qij[st] = np.array([0, 0, 0, 1], dtype=poses.dtype)
    return tij, qij


def _pixel_grid(ht, wd, intr, dtype):
    fx, fy, cx, cy = (dtype(v) for v in intr)
    v, u = np.meshgrid(np.arange(ht, dtype=dtype), np.arange(wd, dtype=dtype), indexing="ij")
    return u.reshape(-1), v.reshape(-1), (u.reshape(-1) - cx) / fx, (v.reshape(-1) - cy) / fy


# --------------------------------------------------------------------------
# a6: per-edge Jacobian / Hessian kernel
# --------------------------------------------------------------------------
def jacobian_edges(poses, disps, intr, targets, weights, ii, jj, dtype=np.float64, chunk=256):
    """Restates projective_transform_kernel, src/droid_kernels.cu:220-468.

    Returns Hs[4,E,6,6] (ii,ij,ji,jj), vs[2,E,6], Eii[E,6,HW], Eij[E,6,HW],
    Cii[E,HW], bz[E,HW].
    """
    poses = np.asarray(poses, dtype)
    N, ht, wd = disps.shape
    HW = ht * wd
    E = len(ii)
    fx, fy, cx, cy = (dtype(v) for v in intr)
    _, _, X0, X1 = _pixel_grid(ht, wd, intr, dtype)
    dflat = np.asarray(disps, dtype).reshape(N, HW)
    tg = np.asarray(targets, dtype).reshape(E, 2, HW)
    wt = np.asarray(weights, dtype).reshape(E, 2, HW)

    Hs = np.zeros((4, E, 6, 6), dtype)
    vs = np.zeros((2, E, 6), dtype)
    Eii = np.zeros((E, 6, HW), dtype)
    Eij = np.zeros((E, 6, HW), dtype)
    Cii = np.zeros((E, HW), dtype)
    bz = np.zeros((E, HW), dtype)

    for s in range(0, E, chunk):
        e = slice(s, min(E, s + chunk))
        i_, j_ = ii[e], jj[e]
        n = len(i_)
        tij, qij = _edge_transforms(poses, i_, j_)
        stereo = (i_ == j_)[:, None]
        d0 = dflat[i_]                                           # [n,HW]
        Xi = np.stack([np.broadcast_to(X0, (n, HW)), np.broadcast_to(X1, (n, HW)),
                       np.ones((n, HW), dtype), d0], axis=-1)    # [n,HW,4]
        Xj = act_se3(tij[:, None, :], qij[:, None, :], Xi)
        x, y, z, h = Xj[..., 0], Xj[..., 1], Xj[..., 2], Xj[..., 3]
        near = z < dtype(MIN_DEPTH)
        with np.errstate(divide="ignore", invalid="ignore"):
            d = np.where(near, dtype(0), dtype(1) / z)
        d2 = d * d
        wu = np.where(near, dtype(0), dtype(.001) * wt[e, 0])
        wv = np.where(near, dtype(0), dtype(.001) * wt[e, 1])
        ru = tg[e, 0] - (fx * d * x + cx)
        rv = tg[e, 1] - (fy * d * y + cy)
        o = np.zeros_like(x)
        Jju = fx * np.stack([h * d, o, -x * h * d2, -x * y * d2, 1 + x * x * d2, -y * d], axis=-1)
        Jjv = fy * np.stack([o, h * d, -y * h * d2, -1 - y * y * d2, x * y * d2, x * d], axis=-1)
        Jzu = fx * (tij[:, None, 0] * d - tij[:, None, 2] * (x * d2))
        Jzv = fy * (tij[:, None, 1] * d - tij[:, None, 2] * (y * d2))
        Cii[e] = wu * Jzu * Jzu + wv * Jzv * Jzv
        bz[e] = wu * ru * Jzu + wv * rv * Jzv
        wu = np.where(stereo, dtype(0), wu)       # src/droid_kernels.cu:367,400
        wv = np.where(stereo, dtype(0), wv)
        Jiu = -adj_se3(tij[:, None, :], qij[:, None, :], Jju)
        Jiv = -adj_se3(tij[:, None, :], qij[:, None, :], Jjv)
        Ju = np.concatenate([Jiu, Jju], axis=-1)   # [n,HW,12]
        Jv = np.concatenate([Jiv, Jjv], axis=-1)
        H = np.einsum("eh,ehn,ehm->enm", wu, Ju, Ju) + np.einsum("eh,ehn,ehm->enm", wv, Jv, Jv)
        Hs[0, e] = H[:, :6, :6]
        Hs[1, e] = H[:, :6, 6:]
        Hs[2, e] = H[:, 6:, :6]
        Hs[3, e] = H[:, 6:, 6:]
        g = np.einsum("eh,ehn->en", wu * ru, Ju) + np.einsum("eh,ehn->en", wv * rv, Jv)
        vs[0, e] = g[:, :6]
        vs[1, e] = g[:, 6:]
        Eii[e] = np.transpose((wu * Jzu)[..., None] * Jiu + (wv * Jzv)[..., None] * Jiv, (0, 2, 1))
        Eij[e] = np.transpose((wu * Jzu)[..., None] * Jju + (wv * Jzv)[..., None] * Jjv, (0, 2, 1))
    return Hs, vs, Eii, Eij, Cii, bz


# --------------------------------------------------------------------------
# A.2 bookkeeping (integer exact)
# --------------------------------------------------------------------------
def build_tables(ii, jj, t0, t1):
    """src/droid_kernels.cu:1416-1424: ts, ii_exp, jj_exp, (kx, kk_exp)=unique."""
    ii = np.asarray(ii, np.int64)
    jj = np.asarray(jj, np.int64)
    ts = np.arange(t0, t1, dtype=np.int64)
    ii_exp = np.concatenate([ts, ii])
    jj_exp = np.concatenate([ts, jj])
    kx, kk_exp = np.unique(ii_exp, return_inverse=True)
    return dict(ts=ts, ii_exp=ii_exp, jj_exp=jj_exp, kx=kx.astype(np.int64),
                kk_exp=kk_exp.astype(np.int64))


def csr_by_source(ii, kx):
    """Edges grouped by source frame in ascending edge order: the segment lists
    accum_cuda builds on the host (src/droid_kernels.cu:993-1031), with the
    unspecified argsort tie order fixed to ascending edge index."""
    ii = np.asarray(ii, np.int64)
    pos = {int(k): n for n, k in enumerate(kx)}
    lists = [[] for _ in kx]
    for e, i in enumerate(ii):
        lists[pos[int(i)]].append(e)
    ptr = np.zeros(len(kx) + 1, np.int64)
    for n, l in enumerate(lists):
        ptr[n + 1] = ptr[n] + len(l)
    idx = np.array([e for l in lists for e in l], np.int64)
    return ptr, idx


def schur_rows(tab, t0, t1):
    """Rows of E that take part in the Schur complement, grouped per pose.
    src/droid_kernels.cu:1316-1328 (inclusive upper bound at :1323; a row with
    pose == t1 indexes past graph[P] in the reference, callers never produce it)."""
    P = t1 - t0
    graph = [[] for _ in range(P)]
    index = [[] for _ in range(P)]
    for n, (j, k) in enumerate(zip(tab["jj_exp"], tab["kk_exp"])):
        if t0 <= j <= t1:
            assert j < t1, "row with pose == t1 is out of bounds in the reference"
            graph[j - t0].append(int(k))
            index[j - t0].append(n)
    return graph, index


def schur_triples(tab, t0, t1):
    """(pose_a, pose_b, row_a, row_b, depth) table in the reference's order.
    src/droid_kernels.cu:1330-1347."""
    graph, index = schur_rows(tab, t0, t1)
    P = t1 - t0
    out = []
    by_depth = []
    for i in range(P):
        d = {}
        for k, n in zip(graph[i], index[i]):
            d.setdefault(k, []).append(n)
        by_depth.append(d)
    for i in range(P):
        for j in range(P):
            for k, n in zip(graph[i], index[i]):
                for m in by_depth[j].get(k, ()):
                    out.append((i, j, n, m, k))
    return np.array(out, np.int64).reshape(-1, 5)


def accum(data, ix, jx):
    """out[j] = sum_{i: ix[i]==jx[j]} data[i].  src/droid_kernels.cu:899-919,993-1043."""
    out = np.zeros((len(jx),) + data.shape[1:], data.dtype)
    pos = {int(k): n for n, k in enumerate(jx)}
    for i, k in enumerate(ix):
        n = pos.get(int(k))
        if n is not None:
            out[n] += data[i]
    return out


# --------------------------------------------------------------------------
# a7-a12: assembly, Schur complement, solve, back-substitution, retraction
# --------------------------------------------------------------------------
def assemble_pose_system(Hs, vs, ii, jj, t0, t1):
    """SparseBlock::update_lhs/update_rhs, src/droid_kernels.cu:1176-1218 (fp64)."""
    P = t1 - t0
    A = np.zeros((P, 6, P, 6), np.float64)
    b = np.zeros((P, 6), np.float64)
    E = len(ii)
    rows = np.concatenate([ii, ii, jj, jj]) - t0
    cols = np.concatenate([ii, jj, ii, jj]) - t0
    blocks = np.asarray(Hs, np.float64).reshape(4 * E, 6, 6)
    for n in range(4 * E):
        i, j = rows[n], cols[n]
        if i >= 0 and j >= 0:
            A[i, :, j, :] += blocks[n]
    vrows = np.concatenate([ii, jj]) - t0
    vb = np.asarray(vs, np.float64).reshape(2 * E, 6)
    for n in range(2 * E):
        if vrows[n] >= 0:
            b[vrows[n]] += vb[n]
    return A.reshape(6 * P, 6 * P), b.reshape(6 * P)


def depth_system(Cii, bz, Eii, Eij, disps, disps_sens, eta, ii, tab, t0, t1, alpha, dtype):
    """C, w, Q, E rows.  src/droid_kernels.cu:1474-1481 (alpha=0.05) / 1872-1879 (alpha=0.001)."""
    kx = tab["kx"]
    K = len(kx)
    HW = Cii.shape[1]
    dk = np.asarray(disps, dtype).reshape(len(disps), HW)[kx]
    sk = np.asarray(disps_sens, dtype).reshape(len(disps_sens), HW)[kx]
    m = (sk > 0).astype(dtype)
    et = np.broadcast_to(np.asarray(eta, dtype).reshape(-1, HW), (K, HW))
    C = accum(Cii, ii, kx) + m * dtype(alpha) + (1 - m) * et
    w = accum(bz, ii, kx) - m * dtype(alpha) * (dk - sk)
    Q = dtype(1.0) / C
    Ei = accum(Eii, ii, tab["ts"])
    E = np.concatenate([Ei, Eij], axis=0)
    return C, w, Q, E


def schur_system(E, Q, w, tab, t0, t1, dtype):
    """S (dense 6P x 6P, fp64 sum of fp-`dtype` blocks) and s_v.
    src/droid_kernels.cu:1046-1138 (kernels), 1297-1391 (assembly)."""
    P = t1 - t0
    S = np.zeros((P, 6, P, 6), np.float64)
    sv = np.zeros((P, 6), np.float64)
    tri = schur_triples(tab, t0, t1)
    for a, b, n, m, k in tri:
        blk = (E[n] * Q[k]) @ E[m].T            # EEt6x6_kernel
        S[a, :, b, :] += blk.astype(dtype).astype(np.float64)
    qw = Q * w
    for n in range(E.shape[0]):                 # Ev6x1_kernel over all rows
        p = tab["jj_exp"][n] - t0
        v = (E[n] @ qw[tab["kk_exp"][n]]).astype(dtype)
        if p >= 0:
            sv[p] += v
    return S.reshape(6 * P, 6 * P), sv.reshape(6 * P)


def damped_solve(A, b, lm, ep):
    """diag += ep + lm*diag; fp64 LLT; dx = 0 when a pivot is <= 0.
    src/droid_kernels.cu:1248-1288, 180-218.  lm/ep arrive as float32."""
    L = np.array(A, np.float64)
    d = np.diag(L).copy()
    L[np.diag_indices_from(L)] = d + (float(np.float32(ep)) + float(np.float32(lm)) * d)
    try:
        c = np.linalg.cholesky(L)
    except np.linalg.LinAlgError:
        return np.zeros(len(b), np.float32), False
    y = np.linalg.solve(c, b)
    x = np.linalg.solve(c.T, y)
    return x.astype(np.float32), True


def block_tridiagonal_cyclic_reduction(D, C, y):
    """Solves the symmetric positive definite block-tridiagonal system with diagonal blocks D[1..m], couplings
    C[a] = A(a+1, a) (rows: unknowns of a+1) and right-hand sides y[1..m] by block cyclic reduction, in the order the
    domain-decomposed banded solve works (csrc/ba_spike.cuh): at level l (h = 2^l) every s with s mod 2h == h is
    eliminated -- L_s = chol(D_s), Xa = A(s-h, s) L_s^-T, Xb = A(s+h, s) L_s^-T, D_{s-h} -= Xa Xa^T,
    D_{s+h} -= Xb Xb^T, A(s+h, s-h) = -Xb Xa^T, same for y -- until one block is left, then back through the levels.
    No counterpart in the reference (it hands the whole reduced camera system to Eigen::SimplicialLLT,
    src/droid_kernels.cu:1248-1269); pinned against a dense solve in tests/test_oracle_cpu.py.
    D, C, y: dicts keyed 1..m / 1..m-1 / 1..m of float64 arrays.  Returns {s: x_s} and the number of factor levels."""
    m = len(D)
    D = {s: np.array(D[s], np.float64) for s in D}
    y = {s: np.array(y[s], np.float64) for s in y}
    coup = {(0, a): np.array(C[a], np.float64) for a in C}
    L, z, Xa, Xb = {}, {}, {}, {}
    h, lev = 1, 0
    while True:
        active = [s for s in range(1, m + 1) if s % h == 0]
        if len(active) == 1:
            last = active[0]
            L[last] = np.linalg.cholesky(D[last])
            z[last] = np.linalg.solve(L[last], y[last])
            break
        for s in [s for s in range(1, m + 1) if s % (2 * h) == h]:
            L[s] = np.linalg.cholesky(D[s])
            z[s] = np.linalg.solve(L[s], y[s])
            a, b = s - h, s + h
            if a >= 1:
                Xa[s] = np.linalg.solve(L[s], coup[(lev, a)]).T          # A(a,s) = A(s,a)^T
                D[a] -= Xa[s] @ Xa[s].T
                y[a] -= Xa[s] @ z[s]
            if b <= m:
                Xb[s] = np.linalg.solve(L[s], coup[(lev, s)].T).T
                D[b] -= Xb[s] @ Xb[s].T
                y[b] -= Xb[s] @ z[s]
            if a >= 1 and b <= m:
                coup[(lev + 1, a)] = -Xb[s] @ Xa[s].T
        h, lev = 2 * h, lev + 1
    x = {last: np.linalg.solve(L[last].T, z[last])}
    levels = lev + 1
    while lev > 0:
        lev, h = lev - 1, h // 2
        for s in [s for s in range(1, m + 1) if s % (2 * h) == h]:
            r = z[s].copy()
            if s - h >= 1:
                r -= Xa[s].T @ x[s - h]
            if s + h <= m:
                r -= Xb[s].T @ x[s + h]
            x[s] = np.linalg.solve(L[s].T, r)
    return x, levels


def back_substitute(E, Q, w, dx, tab, t0, t1, dtype):
    """EvT6x1_kernel + dz.  src/droid_kernels.cu:1140-1160, 1486-1495.
    Rows whose pose index is <= 0 or >= P contribute nothing (:1150)."""
    P = t1 - t0
    dxp = np.asarray(dx, dtype).reshape(P, 6)
    dw = np.zeros((E.shape[0], E.shape[2]), dtype)
    for n in range(E.shape[0]):
        p = tab["jj_exp"][n] - t0
        if 0 < p < P:
            dw[n] = dxp[p] @ E[n]
    return Q * (w - accum(dw, tab["ii_exp"], tab["kx"]))


def retract(poses, disps, dx, dz, tab, t0, t1):
    """pose_retr_kernel / disp_retr_kernel.  src/droid_kernels.cu:943-991."""
    dt = poses.dtype
    if dx is not None:
        t, q = retr_se3(np.asarray(dx, dt).reshape(-1, 6), poses[t0:t1, :3], poses[t0:t1, 3:])
        poses[t0:t1, :3] = t
        poses[t0:t1, 3:] = q
    if dz is not None:
        N = disps.shape[0]
        flat = disps.reshape(N, -1)
        flat[tab["kx"]] += np.asarray(dz, disps.dtype)


def ba(poses, disps, intr, disps_sens, targets, weights, eta, ii, jj, t0, t1,
       iterations, lm, ep, motion_only, dtype=np.float32, alpha=0.05, trace=None):
    """ba_cuda, src/droid_kernels.cu:1394-1512.  Mutates poses/disps, returns (dx, dz)."""
    ii = np.asarray(ii, np.int64)
    jj = np.asarray(jj, np.int64)
    tab = build_tables(ii, jj, t0, t1)
    dx = dz = None
    for _ in range(iterations):
        Hs, vs, Eii, Eij, Cii, bz = jacobian_edges(poses, disps, intr, targets, weights, ii, jj, dtype)
        A, b = assemble_pose_system(Hs, vs, ii, jj, t0, t1)
        if motion_only:
            dx, ok = damped_solve(A, b, lm, ep)
            retract(poses, disps, dx, None, tab, t0, t1)
            if trace is not None:
                trace.append(dict(H=A, v=b, dx=dx.copy(), ok=ok))
            continue
        C, w, Q, E = depth_system(Cii, bz, Eii, Eij, disps, disps_sens, eta, ii, tab, t0, t1, alpha, dtype)
        S, sv = schur_system(E, Q, w, tab, t0, t1, dtype)
        dx, ok = damped_solve(A - S, b - sv, lm, ep)
        dz = back_substitute(E, Q, w, dx, tab, t0, t1, dtype)
        if trace is not None:
            trace.append(dict(H=A - S, v=b - sv, dx=dx.copy(), dz=dz.copy(), ok=ok,
                              Hs=Hs, vs=vs, E=E, Q=Q, w=w, C=C))
        retract(poses, disps, dx, dz, tab, t0, t1)
    return (None if dx is None else dx.reshape(-1, 6)), dz


def ba_extend(poses, disps, intr, disps_sens, targets, weights, eta, ii, jj, H, v, A_prior,
              t0, t1, iterations, lm, ep, motion_only, skip_solve, dtype=np.float32):
    """ba_cuda_extend, src/droid_kernels.cu:1514-1678.  H, v (float64) are overwritten
    with the undamped reduced system; the solve uses L = Ad + Adprior where Adprior
    is a copy of Ad overwritten on its leading block by A_prior (:1624-1630)."""
    ii = np.asarray(ii, np.int64)
    jj = np.asarray(jj, np.int64)
    tab = build_tables(ii, jj, t0, t1)
    dx = dz = None
    for _ in range(iterations):
        Hs, vs, Eii, Eij, Cii, bz = jacobian_edges(poses, disps, intr, targets, weights, ii, jj, dtype)
        A, b = assemble_pose_system(Hs, vs, ii, jj, t0, t1)
        if motion_only:
            dx, _ = damped_solve(A, b, lm, ep)
            retract(poses, disps, dx, None, tab, t0, t1)
            continue
        C, w, Q, E = depth_system(Cii, bz, Eii, Eij, disps, disps_sens, eta, ii, tab, t0, t1, 0.05, dtype)
        S, sv = schur_system(E, Q, w, tab, t0, t1, dtype)
        Ad, vd = A - S, b - sv
        Apr = Ad.copy()
        r, c = A_prior.shape
        Apr[:r, :c] = A_prior
        H[:, :] = Ad[:H.shape[0], :H.shape[1]]
        v[:] = vd[:H.shape[0]]
        if skip_solve:
            return dx, dz
        dx, _ = damped_solve(Ad + Apr, vd, lm, ep)
        dz = back_substitute(E, Q, w, dx, tab, t0, t1, dtype)
        retract(poses, disps, dx, dz, tab, t0, t1)
    return (None if dx is None else dx.reshape(-1, 6)), dz


class BACore:
    """BACore::{init,hessian,optimize,retract}, src/bacore.h:4-70, src/droid_kernels.cu:1786-1956."""

    def init(self, poses, disps, intr, disps_sens, targets, weights, eta, ii, jj,
             t0, t1, iterations, lm, ep, motion_only, dtype=np.float32):
        self.poses, self.disps, self.intr, self.disps_sens = poses, disps, intr, disps_sens
        self.targets, self.weights, self.eta = targets, weights, eta
        self.ii, self.jj = np.asarray(ii, np.int64), np.asarray(jj, np.int64)
        self.t0, self.t1, self.lm, self.ep, self.dtype = t0, t1, lm, ep, dtype
        self.tab = build_tables(self.ii, self.jj, t0, t1)

    def hessian(self, H, v):
        Hs, vs, Eii, Eij, Cii, bz = jacobian_edges(self.poses, self.disps, self.intr, self.targets,
                                                   self.weights, self.ii, self.jj, self.dtype)
        A, b = assemble_pose_system(Hs, vs, self.ii, self.jj, self.t0, self.t1)
        self.C, self.w, self.Q, self.E = depth_system(
            Cii, bz, Eii, Eij, self.disps, self.disps_sens, self.eta, self.ii, self.tab,
            self.t0, self.t1, 0.001, self.dtype)
        S, sv = schur_system(self.E, self.Q, self.w, self.tab, self.t0, self.t1, self.dtype)
        H[:, :] = (A - S)[:H.shape[0], :H.shape[1]]
        v[:] = (b - sv)[:H.shape[0]]

    def optimize(self, H, v):
        self.dx, _ = damped_solve(H, v, self.lm, self.ep)

    def retract(self, dx):
        dx32 = np.asarray(dx, np.float64).astype(np.float32)
        dz = back_substitute(self.E, self.Q, self.w, dx32, self.tab, self.t0, self.t1, self.dtype)
        retract(self.poses, self.disps, dx32, dz, self.tab, self.t0, self.t1)
        return dx32.reshape(-1, 6), dz


# --------------------------------------------------------------------------
# a16-a18: projmap, frame_distance, iproj, depth_filter
# --------------------------------------------------------------------------
def projmap(poses, disps, intr, ii, jj, dtype=np.float32):
    """projmap_kernel, src/droid_kernels.cu:471-560 (no stereo special case)."""
    poses = np.asarray(poses, dtype)
    N, ht, wd = disps.shape
    HW = ht * wd
    fx, fy, cx, cy = (dtype(v) for v in intr)
    u, v, X0, X1 = _pixel_grid(ht, wd, intr, dtype)
    tij, qij = _edge_transforms(poses, ii, jj, stereo_special=False)
    n = len(ii)
    d0 = np.asarray(disps, dtype).reshape(N, HW)[ii]
    Xi = np.stack([np.broadcast_to(X0, (n, HW)), np.broadcast_to(X1, (n, HW)), np.ones((n, HW), dtype), d0], -1)
    Xj = act_se3(tij[:, None], qij[:, None], Xi)
    z = Xj[..., 2]
    coords = np.zeros((n, HW, 3), dtype)
    ok = z > dtype(0.01)
    with np.errstate(divide="ignore", invalid="ignore"):
        coords[..., 0] = np.where(ok, fx * (Xj[..., 0] / z) + cx, u)
        coords[..., 1] = np.where(ok, fy * (Xj[..., 1] / z) + cy, v)
    valid = (z > dtype(MIN_DEPTH)).astype(dtype)
    return coords.reshape(n, ht, wd, 3), valid.reshape(n, ht, wd, 1)


def ba2gtsam(H, v, A):
    """BA2GTSAM, dbaf/depth_video.py:20-29 with the 6x6 matrix A given (the reference builds it from
    gtsam.Pose3: the rows of -Tbc.inverse().AdjointMap() swapped, :21-23): J = blockdiag(A),
    returns J^T H J and J^T v."""
    H, v, A = np.asarray(H, np.float64), np.asarray(v, np.float64), np.asarray(A, np.float64)
    ss = H.shape[0] // 6
    J = np.zeros_like(H)
    for i in range(ss):
        J[i * 6:i * 6 + 6, i * 6:i * 6 + 6] = A
    return J.T @ H @ J, J.T @ v


def reproject(poses, disps, intr, ii, jj, dtype=np.float32):
    """DepthVideo.reproject, dbaf/depth_video.py:221-229 -> projective_transform without Jacobians,
    dbaf/geom/projective_ops.py:96-125: iproj with the intrinsics of frame ii (:18-31), Gij = Tj * Ti^-1 with
    the stereo special case (:104-106), act (:67-69), proj with Z < 0.5*MIN_DEPTH -> 1 (:38-46),
    valid = X1.z > MIN_DEPTH(0.2) & X0.z(=1) > MIN_DEPTH (:114-115).  intr: [4] or [N,4].
    Returns coords [E,ht,wd,2], valid [E,ht,wd,1]."""
    poses = np.asarray(poses, dtype)
    N, ht, wd = disps.shape
    HW = ht * wd
    K = np.asarray(intr, dtype)
    K = np.broadcast_to(K, (N, 4)) if K.ndim == 1 else K
    ii, jj = np.asarray(ii), np.asarray(jj)
    n = len(ii)
    v, u = np.meshgrid(np.arange(ht, dtype=dtype), np.arange(wd, dtype=dtype), indexing="ij")
    u, v = u.reshape(1, HW), v.reshape(1, HW)
    Ki, Kj = K[ii], K[jj]
    X0 = (u - Ki[:, 2:3]) / Ki[:, 0:1]
    X1 = (v - Ki[:, 3:4]) / Ki[:, 1:2]
    tij, qij = _edge_transforms(poses, ii, jj, stereo_special=True)
    d0 = np.asarray(disps, dtype).reshape(N, HW)[ii]
    Xi = np.stack([X0, X1, np.ones((n, HW), dtype), d0], -1)
    Xj = act_se3(tij[:, None], qij[:, None], Xi)
    z = Xj[..., 2]
    zc = np.where(z < dtype(0.1), dtype(1.0), z)
    d = dtype(1.0) / zc
    coords = np.stack([Kj[:, 0:1] * (Xj[..., 0] * d) + Kj[:, 2:3], Kj[:, 1:2] * (Xj[..., 1] * d) + Kj[:, 3:4]], -1)
    valid = (z > dtype(0.2)).astype(dtype)
    return coords.reshape(n, ht, wd, 2).astype(dtype), valid.reshape(n, ht, wd, 1)


def frame_distance(poses, disps, intr, ii, jj, beta, dtype=np.float32):
    """frame_distance_kernel, src/droid_kernels.cu:562-702 (one direction only, n<1 loop)."""
    poses = np.asarray(poses, dtype)
    N, ht, wd = disps.shape
    HW = ht * wd
    fx, fy, cx, cy = (dtype(v) for v in intr)
    beta = dtype(beta)
    u, v, X0, X1 = _pixel_grid(ht, wd, intr, dtype)
    tij, qij = _edge_transforms(poses, ii, jj, stereo_special=False)
    n = len(ii)
    d0 = np.asarray(disps, dtype).reshape(N, HW)[ii]
    Xi = np.stack([np.broadcast_to(X0, (n, HW)), np.broadcast_to(X1, (n, HW)), np.ones((n, HW), dtype), d0], -1)
    out = np.zeros(n, dtype)
    with np.errstate(divide="ignore", invalid="ignore"):
        Xj = act_se3(tij[:, None], qij[:, None], Xi)
        du = fx * (Xj[..., 0] / Xj[..., 2]) + cx - u
        dv = fy * (Xj[..., 1] / Xj[..., 2]) + cy - v
        d1 = np.sqrt(du * du + dv * dv)
        ok1 = Xj[..., 2] > dtype(MIN_DEPTH)
        Yj = Xi[..., :3] + Xi[..., 3:4] * tij[:, None]
        du = fx * (Yj[..., 0] / Yj[..., 2]) + cx - u
        dv = fy * (Yj[..., 1] / Yj[..., 2]) + cy - v
        d2 = np.sqrt(du * du + dv * dv)
        ok2 = Yj[..., 2] > dtype(MIN_DEPTH)
    total = np.full(n, HW * (beta + (1 - beta)), np.float64)
    acc = np.where(ok1, beta * d1, 0).sum(-1, dtype=np.float64) + np.where(ok2, (1 - beta) * d2, 0).sum(-1, dtype=np.float64)
    val = np.where(ok1, beta, 0).sum(-1, dtype=np.float64) + np.where(ok2, 1 - beta, 0).sum(-1, dtype=np.float64)
    with np.errstate(divide="ignore", invalid="ignore"):
        out = np.where(val / (total + 1e-8) < 0.75, 1000.0, acc / val)
    return out.astype(dtype)


def iproj(poses, disps, intr, dtype=np.float32):
    """iproj_kernel, src/droid_kernels.cu:824-895: X = T * (x,y,1,d); out = X[:3]/X[3]."""
    poses = np.asarray(poses, dtype)
    N, ht, wd = disps.shape
    HW = ht * wd
    _, _, X0, X1 = _pixel_grid(ht, wd, intr, dtype)
    Xi = np.stack([np.broadcast_to(X0, (N, HW)), np.broadcast_to(X1, (N, HW)), np.ones((N, HW), dtype),
                   np.asarray(disps, dtype).reshape(N, HW)], -1)
    Xj = act_se3(poses[:N, None, :3], poses[:N, None, 3:], Xi)
    with np.errstate(divide="ignore", invalid="ignore"):
        pts = Xj[..., :3] / Xj[..., 3:4]
    return pts.reshape(N, ht, wd, 3)


def depth_filter(poses, disps, intr, ix, thresh, dtype=np.float32):
    """depth_filter_kernel, src/droid_kernels.cu:706-820.  Neighbours ix-1..ix-3, ix+3..ix+5 (:740)."""
    poses = np.asarray(poses, dtype)
    N, ht, wd = disps.shape
    HW = ht * wd
    fx, fy, cx, cy = (dtype(v) for v in intr)
    _, _, X0, X1 = _pixel_grid(ht, wd, intr, dtype)
    dflat = np.asarray(disps, dtype).reshape(N, HW)
    count = np.zeros((len(ix), HW), dtype)
    for b, i in enumerate(ix):
        t = dtype(thresh[b])
        for nb in range(6):
            j = i - nb - 1 if nb < 3 else i + nb
            if j < 0 or j >= N:
                continue
            tij, qij = rel_se3(poses[i, :3], poses[i, 3:], poses[j, :3], poses[j, 3:])
            Xi = np.stack([X0, X1, np.ones(HW, dtype), dflat[i]], -1)
            Xj = act_se3(tij, qij, Xi)
            with np.errstate(divide="ignore", invalid="ignore"):
                uj = fx * (Xj[:, 0] / Xj[:, 2]) + cx
                vj = fy * (Xj[:, 1] / Xj[:, 2]) + cy
                dj = Xj[:, 3] / Xj[:, 2]
                fu, fv = np.floor(uj), np.floor(vj)
                ok = (fu >= 0) & (fv >= 0) & (fu < wd - 1) & (fv < ht - 1)     # NaN compares false
                u0 = np.where(ok, fu, 0).astype(np.int64)
                v0 = np.where(ok, fv, 0).astype(np.int64)
                dj_img = dflat[j].reshape(ht, wd)
                hit = np.zeros(HW, bool)
                for dv_, du_ in ((0, 0), (0, 1), (1, 0), (1, 1)):
                    dn = dj_img[np.minimum(v0 + dv_, ht - 1), np.minimum(u0 + du_, wd - 1)]
                    hit |= np.abs(1.0 / dj.astype(np.float64) - 1.0 / dn.astype(np.float64)) < float(t)
            count[b] += (ok & hit).astype(dtype)
    return count.reshape(len(ix), ht, wd)


# --------------------------------------------------------------------------
# a1-a5: correlation volume, pyramid, lookup, altcorr
# --------------------------------------------------------------------------
def corr_volume(fmap1, fmap2, out_dtype=np.float16):
    """CorrBlock.corr, dbaf/modules/corr.py:63-71: (f1/4)^T (f2/4), products and sums in
    fp32 (cuBLAS HGEMM with fp32 accumulate), one rounding to `out_dtype`.
    fmap1/fmap2: [n, C, ht, wd]."""
    n, C, ht, wd = fmap1.shape
    a = (np.asarray(fmap1).astype(fmap1.dtype) / fmap1.dtype.type(4.0)).reshape(n, C, ht * wd).astype(np.float32)
    b = (np.asarray(fmap2).astype(fmap2.dtype) / fmap2.dtype.type(4.0)).reshape(n, C, -1).astype(np.float32)
    h2, w2 = fmap2.shape[2:]
    vol = np.einsum("ncp,ncq->npq", a, b, optimize=True)
    return vol.astype(out_dtype).reshape(n, ht, wd, h2, w2)


def avg_pool2(vol):
    """F.avg_pool2d(corr, 2, stride=2) over the last two dims (floor), accumulate in
    fp32 in (y0x0, y0x1, y1x0, y1x1) order, divide by 4, round to the storage type.
    dbaf/modules/corr.py:35-38."""
    h2, w2 = vol.shape[-2:]
    h, w = h2 // 2, w2 // 2
    v = vol[..., :2 * h, :2 * w].astype(np.float32)
    s = ((v[..., 0::2, 0::2] + v[..., 0::2, 1::2]) + v[..., 1::2, 0::2]) + v[..., 1::2, 1::2]
    return (s / np.float32(4.0)).astype(vol.dtype)


def corr_pyramid(fmap1, fmap2, levels=4, out_dtype=np.float16):
    """CorrBlock.__init__, dbaf/modules/corr.py:24-38."""
    pyr = [corr_volume(fmap1, fmap2, out_dtype)]
    for _ in range(levels - 1):
        pyr.append(avg_pool2(pyr[-1]))
    return pyr


def corr_index_forward(volume, coords, r):
    """corr_index_forward_kernel, src/correlation_kernels.cu:19-70.

    For fp16 volumes this reproduces the reference's rounding sequence exactly
    (weights rounded to fp16, every product and every add rounded to fp16, terms
    added in tap order i-outer/j-inner).  For fp32/fp64 the
    multiply-add is done as a fused multiply-add would (computed in float64 then
    rounded), which is how nvcc contracts `corr += s * w`."""
    n, h1, w1, h2, w2 = volume.shape
    T = volume.dtype.type
    rd = 2 * r + 1
    x0 = np.asarray(coords[:, 0], np.float32)
    y0 = np.asarray(coords[:, 1], np.float32)
    fx0 = np.floor(x0)
    fy0 = np.floor(y0)
    dx = (x0 - fx0).astype(np.float32)
    dy = (y0 - fy0).astype(np.float32)
    one = np.float32(1.0)
    w11 = (dx * dy).astype(T)                 # goes to out[i-1][j-1]
    w10 = (dx * (one - dy)).astype(T)         # out[i-1][j]
    w01 = ((one - dx) * dy).astype(T)         # out[i][j-1]
    w00 = ((one - dx) * (one - dy)).astype(T)  # out[i][j]
    out = np.zeros((n, rd, rd, h1, w1), T)
    nn, yy, xx = np.meshgrid(np.arange(n), np.arange(h1), np.arange(w1), indexing="ij")
    finite = np.isfinite(x0) & np.isfinite(y0)
    bx = np.where(finite, fx0, -1e6).astype(np.int64)
    by = np.where(finite, fy0, -1e6).astype(np.int64)

    def madd(acc, s, w, mask):
        if T == np.float16:
            prod = (s.astype(np.float32) * w.astype(np.float32)).astype(np.float16)
            new = (acc.astype(np.float32) + prod.astype(np.float32)).astype(np.float16)
        else:
            new = (acc.astype(np.float64) + s.astype(np.float64) * w.astype(np.float64)).astype(T)
        return np.where(mask, new, acc)

    for i in range(rd + 1):
        for j in range(rd + 1):
            x1 = bx - r + i
            y1 = by - r + j
            inb = (x1 >= 0) & (x1 < w2) & (y1 >= 0) & (y1 < h2)
            s = volume[nn, yy, xx, np.clip(y1, 0, h2 - 1), np.clip(x1, 0, w2 - 1)]
            if i > 0 and j > 0:
                out[:, i - 1, j - 1] = madd(out[:, i - 1, j - 1], s, w11, inb)
            if i > 0 and j < rd:
                out[:, i - 1, j] = madd(out[:, i - 1, j], s, w10, inb)
            if i < rd and j > 0:
                out[:, i, j - 1] = madd(out[:, i, j - 1], s, w01, inb)
            if i < rd and j < rd:
                out[:, i, j] = madd(out[:, i, j], s, w00, inb)
    return out


def corr_index_backward(volume_shape, coords, corr_grad, r):
    """corr_index_backward_kernel, src/correlation_kernels.cu:73-124 (g summed in tap
    order in the storage type, then added to the zero-initialised gradient)."""
    n, h1, w1, h2, w2 = volume_shape
    T = corr_grad.dtype.type
    rd = 2 * r + 1
    x0 = np.asarray(coords[:, 0], np.float32)
    y0 = np.asarray(coords[:, 1], np.float32)
    fx0, fy0 = np.floor(x0), np.floor(y0)
    dx = (x0 - fx0).astype(np.float32)
    dy = (y0 - fy0).astype(np.float32)
    one = np.float32(1.0)
    w11 = (dx * dy).astype(T)
    w10 = (dx * (one - dy)).astype(T)
    w01 = ((one - dx) * dy).astype(T)
    w00 = ((one - dx) * (one - dy)).astype(T)
    grad = np.zeros(volume_shape, T)
    nn, yy, xx = np.meshgrid(np.arange(n), np.arange(h1), np.arange(w1), indexing="ij")
    finite = np.isfinite(x0) & np.isfinite(y0)
    bx = np.where(finite, fx0, -1e6).astype(np.int64)
    by = np.where(finite, fy0, -1e6).astype(np.int64)

    def madd(acc, g, w):
        if T == np.float16:
            prod = (g.astype(np.float32) * w.astype(np.float32)).astype(np.float16)
            return (acc.astype(np.float32) + prod.astype(np.float32)).astype(np.float16)
        return (acc.astype(np.float64) + g.astype(np.float64) * w.astype(np.float64)).astype(T)

    for i in range(rd + 1):
        for j in range(rd + 1):
            x1 = bx - r + i
            y1 = by - r + j
            inb = (x1 >= 0) & (x1 < w2) & (y1 >= 0) & (y1 < h2)
            g = np.zeros((n, h1, w1), T)
            if i > 0 and j > 0:
                g = madd(g, corr_grad[:, i - 1, j - 1], w11)
            if i > 0 and j < rd:
                g = madd(g, corr_grad[:, i - 1, j], w10)
            if i < rd and j > 0:
                g = madd(g, corr_grad[:, i, j - 1], w01)
            if i < rd and j < rd:
                g = madd(g, corr_grad[:, i, j], w00)
            sel = inb
            # each (n,y,x) owns its own [h2,w2] slab, so no two taps collide
            grad[nn[sel], yy[sel], xx[sel], y1[sel], x1[sel]] += g[sel]
    return grad


def altcorr_forward(fmap1, fmap2, coords, r, dtype=np.float32):
    """altcorr_forward_kernel, src/altcorr_kernel.cu:27-149.
    fmap1 [B,H1,W1,C], fmap2 [B,H2,W2,C], coords [B,N,H1,W1,2] (x,y) -> [B,N,rd*rd,H1,W1]
    with channel = iy + rd*ix (:102-105)."""
    B, H1, W1, C = fmap1.shape
    _, H2, W2, _ = fmap2.shape
    N = coords.shape[1]
    rd = 2 * r + 1
    f1 = np.asarray(fmap1, dtype)
    f2 = np.asarray(fmap2, dtype)
    out = np.zeros((B, N, rd * rd, H1, W1), dtype)
    bb, hh, ww = np.meshgrid(np.arange(B), np.arange(H1), np.arange(W1), indexing="ij")
    for n in range(N):
        x = np.asarray(coords[:, n, ..., 0], np.float32)
        y = np.asarray(coords[:, n, ..., 1], np.float32)
        fx0, fy0 = np.floor(x), np.floor(y)
        dx = (x - fx0).astype(dtype)
        dy = (y - fy0).astype(dtype)
        finite = np.isfinite(x) & np.isfinite(y)
        bx = np.where(finite, fx0, -1e6).astype(np.int64)
        by = np.where(finite, fy0, -1e6).astype(np.int64)
        for iy in range(rd + 1):
            for ix in range(rd + 1):
                h2 = by - r + iy
                w2 = bx - r + ix
                inb = (h2 >= 0) & (h2 < H2) & (w2 >= 0) & (w2 < W2)
                g = f2[bb, np.clip(h2, 0, H2 - 1), np.clip(w2, 0, W2 - 1)]
                s = np.where(inb, np.einsum("bhwc,bhwc->bhw", f1, g), 0).astype(dtype)
                if iy > 0 and ix > 0:
                    out[:, n, (iy - 1) + rd * (ix - 1)] += s * (dy * dx)
                if iy > 0 and ix < rd:
                    out[:, n, (iy - 1) + rd * ix] += s * (dy * (1 - dx))
                if iy < rd and ix > 0:
                    out[:, n, iy + rd * (ix - 1)] += s * ((1 - dy) * dx)
                if iy < rd and ix < rd:
                    out[:, n, iy + rd * ix] += s * ((1 - dy) * (1 - dx))
    return out


def altcorr_backward(fmap1, fmap2, coords, corr_grad, r, dtype=np.float32):
    """altcorr_backward_kernel, src/altcorr_kernel.cu:152-286: adjoint of altcorr_forward
    with respect to fmap1 and fmap2; coords_grad is returned as zeros (:335)."""
    B, H1, W1, C = fmap1.shape
    _, H2, W2, _ = fmap2.shape
    N = coords.shape[1]
    rd = 2 * r + 1
    f1 = np.asarray(fmap1, dtype)
    f2 = np.asarray(fmap2, dtype)
    g1 = np.zeros_like(f1)
    g2 = np.zeros_like(f2)
    cg = np.asarray(corr_grad, dtype)
    bb, hh, ww = np.meshgrid(np.arange(B), np.arange(H1), np.arange(W1), indexing="ij")
    for n in range(N):
        x = np.asarray(coords[:, n, ..., 0], np.float32)
        y = np.asarray(coords[:, n, ..., 1], np.float32)
        fx0, fy0 = np.floor(x), np.floor(y)
        dx = (x - fx0).astype(dtype)
        dy = (y - fy0).astype(dtype)
        finite = np.isfinite(x) & np.isfinite(y)
        bx = np.where(finite, fx0, -1e6).astype(np.int64)
        by = np.where(finite, fy0, -1e6).astype(np.int64)
        for iy in range(rd + 1):
            for ix in range(rd + 1):
                h2 = by - r + iy
                w2 = bx - r + ix
                inb = (h2 >= 0) & (h2 < H2) & (w2 >= 0) & (w2 < W2)
                g = np.zeros((B, H1, W1), dtype)
                if iy > 0 and ix > 0:
                    g += cg[:, n, (iy - 1) + rd * (ix - 1)] * dy * dx
                if iy > 0 and ix < rd:
                    g += cg[:, n, (iy - 1) + rd * ix] * dy * (1 - dx)
                if iy < rd and ix > 0:
                    g += cg[:, n, iy + rd * (ix - 1)] * (1 - dy) * dx
                if iy < rd and ix < rd:
                    g += cg[:, n, iy + rd * ix] * (1 - dy) * (1 - dx)
                g = np.where(inb, g, 0)
                f2g = f2[bb, np.clip(h2, 0, H2 - 1), np.clip(w2, 0, W2 - 1)]
                g1 += g[..., None] * f2g
                np.add.at(g2, (bb[inb], h2[inb], w2[inb]), g[inb][:, None] * f1[inb])
    return g1, g2, np.zeros(coords.shape, dtype)


# --------------------------------------------------------------------------
# frame-graph edge selection (dbaf/covisible_graph.py)
# --------------------------------------------------------------------------
def proximity_candidates(t0, t1, t, skip_edge=(), frontend_window=0):
    """Candidate pairs of add_proximity_factors, dbaf/covisible_graph.py:360-377: the (t-t0) x (t-t1) grid in
    row-major order (torch.meshgrid, 'ij'), then -- when skip edges are configured and the grid's source range
    spans exactly frontend_window frames -- one pair (newest frame, oldest + s) per skip offset s with
    oldest + s > 0.  Returns ii, jj (int64) and cc, the size of the grid."""
    ix = np.arange(t0, t, dtype=np.int64)
    jx = np.arange(t1, t, dtype=np.int64)
    ii = np.repeat(ix, len(jx))
    jj = np.tile(jx, len(ix))
    cc = len(ii)
    if len(skip_edge) and cc and ii.max() - ii.min() == frontend_window - 1:
        jj_add = ii.min() + np.asarray(skip_edge, np.int64)
        jj_add = jj_add[jj_add > 0]
        ii = np.concatenate([ii, np.full(len(jj_add), ii.max(), np.int64)])
        jj = np.concatenate([jj, jj_add])
    return ii, jj, cc


def proximity_edges(d, ii, jj, cc, ex_ii, ex_jj, t0, t1, t, rad, nms, thresh, stereo, max_factors):
    """The edge list add_proximity_factors hands to add_factors, dbaf/covisible_graph.py:379-439, given the
    candidates (ii, jj, cc of proximity_candidates), their float32 distances d and the existing edges
    ex_ii/ex_jj (= cat(ii, ii_bad, ii_inac) of the graph, :383-384).  Returns int64 arrays (es_i, es_j)."""
    d = np.array(d, np.float32, copy=True)
    W = t - t1
    d[ii - rad < jj] = np.inf                                   # :380
    d[d > 100] = np.inf                                         # :381

    def stamp(i, j):                                            # :386-394 and :425-433
        lim = max(min(abs(i - j) - 2, nms), 0)
        for di in range(-nms, nms + 1):
            for dj in range(-nms, nms + 1):
                if abs(di) + abs(dj) <= lim:
                    i1, j1 = i + di, j + dj
                    if t0 <= i1 < t and t1 <= j1 < t:
                        d[(i1 - t0) * W + (j1 - t1)] = np.inf

    for i, j in zip(np.asarray(ex_ii).tolist(), np.asarray(ex_jj).tolist()):
        stamp(i, j)
    es = []
    for i in range(t0, t):                                      # :396-406
        if stereo:
            es.append((i, i))
            d[(i - t0) * W + (i - t1)] = np.inf
        for j in range(max(i - rad - 1, 0), i):
            es.append((i, j))
            es.append((j, i))
            if (i - t0) * W + (j - t1) >= 0:
                d[(i - t0) * W + (j - t1)] = np.inf
    order = np.argsort(d, kind="stable")                        # :408 (ties: lowest index first)
    for k in order.tolist():                                    # :409-433
        if k >= cc:
            continue
        if d[k] > thresh:
            continue
        if len(es) > max_factors:
            break
        i, j = int(ii[k]), int(jj[k])
        es.append((i, j))
        es.append((j, i))
        stamp(i, j)
    if len(ii) > cc:                                            # :435-439
        k = cc + int(np.argsort(d[cc:len(ii)], kind="stable")[0])   # NaN sorts last, as in torch.argsort
        if d[k] < thresh and d[k] > 0:
            es.append((int(ii[k]), int(jj[k])))
            es.append((int(jj[k]), int(ii[k])))
    es = np.asarray(es, np.int64).reshape(-1, 2)
    return es[:, 0].copy(), es[:, 1].copy()


def filter_repeated_edges(ii, jj, ex_ii, ex_jj):
    """CovisibleGraph.__filter_repeated_edges, dbaf/covisible_graph.py:61-72: drop the new edges that the graph
    already holds (active or inactive); order kept, duplicates inside the new list are not removed."""
    have = set(zip(np.asarray(ex_ii).tolist(), np.asarray(ex_jj).tolist()))
    keep = np.array([(int(i), int(j)) not in have for i, j in zip(ii, jj)], bool).reshape(-1)
    return np.asarray(ii)[keep], np.asarray(jj)[keep]
