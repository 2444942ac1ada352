"""Seeded synthetic keyframe windows shaped like DBA-Fusion's frontend and backend inputs.

The shapes, intrinsics and graph rules follow SURVEY.md section 8(d): TUM-VI (64x64 and the
real 55x55), KITTI-360 (28x107) and the 1000-keyframe global graph.  Everything is numpy
so the same arrays feed the CUDA path (through torch) and the CPU oracle.

Reference shapes: dbaf/depth_video.py:50-66 (buffers), dbaf/covisible_graph.py:214-342
(targets/weights/eta handed to `ba`), demo_vio_tumvi.py:55-65 (image resize + /8).
"""
from __future__ import annotations

import numpy as np

CONFIGS = {
    # name: (ht, wd, P keyframes, E edges, (fx, fy, cx, cy))
    "c1": (64, 64, 8, 24, (23.87, 23.87, 31.87, 32.11)),
    "c2": (64, 64, 12, 48, (23.87, 23.87, 31.87, 32.11)),
    "c2_real": (55, 55, 12, 48, (20.66, 20.66, 27.57, 27.78)),
    "c3": (28, 107, 12, 48, (60.07, 59.87, 52.34, 24.19)),
    "c4": (64, 64, 12, 36, (23.87, 23.87, 31.87, 32.11)),
    "c5": (64, 64, 1000, 20000, (23.87, 23.87, 31.87, 32.11)),
    "tiny": (16, 24, 5, 10, (12.0, 12.5, 11.5, 7.5)),
    "g24": (64, 64, 24, 480, (23.87, 23.87, 31.87, 32.11)),  # dense banded graph at the full 64x64 frame size
    "g40": (24, 32, 40, 160, (14.0, 14.5, 15.5, 11.5)),     # small banded global graph (6P = 240 > 160)
}


# fixed per-config seed slots: the golden fixtures depend on them, so new configs append
_SEED_SLOT = {"c1": 0, "c2": 1, "c2_real": 2, "c3": 3, "c4": 4, "c5": 5, "tiny": 6, "g40": 7, "g24": 8}


def _qmul(a, b):
    ax, ay, az, aw = a[..., 0], a[..., 1], a[..., 2], a[..., 3]
    bx, by, bz, bw = b[..., 0], b[..., 1], b[..., 2], b[..., 3]
    return np.stack([aw * bx + ax * bw + ay * bz - az * by,
                     aw * by - ax * bz + ay * bw + az * bx,
                     aw * bz + ax * by - ay * bx + az * bw,
                     aw * bw - ax * bx - ay * by - az * bz], -1)


def _qrot(q, v):
    qv = q[..., :3]
    uv = 2.0 * np.cross(qv, v)
    return v + q[..., 3:4] * uv + np.cross(qv, uv)


def _qexp(phi):
    th = np.linalg.norm(phi, axis=-1, keepdims=True)
    s = np.where(th < 1e-8, 0.5, np.sin(0.5 * np.maximum(th, 1e-12)) / np.maximum(th, 1e-12))
    return np.concatenate([s * phi, np.cos(0.5 * th)], -1)


def _smooth(rng, shape, lo, hi, passes=3):
    f = rng.uniform(lo, hi, shape)
    for _ in range(passes):
        f = (f + np.roll(f, 1, -1) + np.roll(f, -1, -1) + np.roll(f, 1, -2) + np.roll(f, -1, -2)) / 5.0
    return f


def trajectory(rng, n):
    """World->camera poses [n,7] (tx,ty,tz,qx,qy,qz,qw) of a smooth forward motion."""
    t_wc = np.zeros((n, 3))
    q_wc = np.zeros((n, 4))
    q_wc[:, 3] = 1
    for k in range(1, n):
        dq = _qexp(rng.normal(0, 0.01, 3))
        q_wc[k] = _qmul(q_wc[k - 1], dq)
        step = rng.normal(0, 0.05, 3) + np.array([0, 0, 0.05])
        t_wc[k] = t_wc[k - 1] + _qrot(q_wc[k - 1], step)
    q_cw = q_wc * np.array([-1, -1, -1, 1.0])
    t_cw = -_qrot(q_cw, t_wc)
    return np.concatenate([t_cw, q_cw], -1)


def frame_graph(rng, n_frames, n_edges, t0=1, radius=2, max_span=None):
    """Bidirectional temporal neighbours within `radius`, then skip edges up to n_edges
    (the shape add_proximity_factors produces, dbaf/covisible_graph.py:357-441)."""
    pairs = []
    for i in range(n_frames):
        for d in range(1, radius + 1):
            if i + d < n_frames:
                pairs.append((i, i + d))
                pairs.append((i + d, i))
    pairs = [p for p in pairs if max(p) >= t0]
    have = set(pairs)
    span = max_span or max(3, n_frames // 2)
    guard = 0
    while len(pairs) < n_edges and guard < 100 * n_edges:
        guard += 1
        i = int(rng.integers(0, n_frames))
        d = int(rng.integers(radius + 1, span + 1))
        j = i + d if rng.random() < 0.5 else i - d
        if 0 <= j < n_frames and (i, j) not in have and max(i, j) >= t0:
            have.add((i, j))
            pairs.append((i, j))
    pairs = pairs[:n_edges]
    ii = np.array([p[0] for p in pairs], np.int64)
    jj = np.array([p[1] for p in pairs], np.int64)
    return ii, jj


def banded_graph(n_frames, reach=10):
    """C5: every frame has out-edges to +-1..+-reach (about 2*reach*n edges)."""
    ii, jj = [], []
    for i in range(n_frames):
        for d in range(1, reach + 1):
            for j in (i - d, i + d):
                if 0 <= j < n_frames:
                    ii.append(i)
                    jj.append(j)
    return np.array(ii, np.int64), np.array(jj, np.int64)


def reproject(poses, disps, intr, ii, jj):
    """Pixel coordinates of frame-ii pixels seen from frame jj (float64), plus depth Z."""
    fx, fy, cx, cy = intr
    n, ht, wd = disps.shape
    v, u = np.meshgrid(np.arange(ht, dtype=np.float64), np.arange(wd, dtype=np.float64), indexing="ij")
    X = np.stack([(u - cx) / fx, (v - cy) / fy, np.ones_like(u)], -1)[None]         # [1,ht,wd,3]
    ti, qi = poses[ii, :3], poses[ii, 3:]
    tj, qj = poses[jj, :3], poses[jj, 3:]
    qi_inv = qi * np.array([-1, -1, -1, 1.0])
    qij = _qmul(qj, qi_inv)
    tij = tj - _qrot(qij, ti)
    Y = _qrot(qij[:, None, None, :], X) + disps[ii][..., None] * tij[:, None, None, :]
    z = Y[..., 2]
    zs = np.where(np.abs(z) < 1e-6, 1e-6, z)
    return np.stack([fx * Y[..., 0] / zs + cx, fy * Y[..., 1] / zs + cy], 1), z   # [E,2,ht,wd]


def make_window(config="c2", seed=0, t0=1, sensor_frac=0.0, near_frac=0.0, stereo_edges=0,
                pose_noise=0.01, n_extra_frames=0, heuristics=False, graph="frontend", reach=None):
    """Returns a dict of float32/int64 numpy arrays: poses [N,7], disps, disps_sens [N,ht,wd],
    intrinsics [4], targets, weights [E,2,ht,wd], eta [K,ht,wd], ii, jj [E], t0, t1, plus the
    ground truth poses_gt / disps_gt."""
    ht, wd, P, E, intr = CONFIGS[config]
    rng = np.random.default_rng(1234 + seed + 1000 * _SEED_SLOT[config])
    n = P + n_extra_frames
    poses_gt = trajectory(rng, n)
    disps_gt = _smooth(rng, (n, ht, wd), 0.3, 2.0)
    if graph == "banded":
        ii, jj = banded_graph(n, reach=reach or max(1, E // (2 * n)))
    else:
        ii, jj = frame_graph(rng, n, E, t0=t0)
    if stereo_edges:
        s = rng.integers(t0, n, stereo_edges)
        ii = np.concatenate([ii, s])
        jj = np.concatenate([jj, s])
    # current estimate = GT perturbed on the left by a small twist; frame 0 stays fixed
    poses = poses_gt.copy()
    for k in range(1, n):
        xi = rng.normal(0, pose_noise, 6)
        dq = _qexp(xi[3:])
        poses[k, 3:] = _qmul(dq, poses_gt[k, 3:])
        poses[k, :3] = _qrot(dq, poses_gt[k, :3]) + xi[:3]
    disps = disps_gt * np.exp(rng.normal(0, 0.05, disps_gt.shape))
    if near_frac > 0:
        # push a fraction of pixels so close that they land behind / near the target camera
        mask = rng.random(disps.shape) < near_frac
        disps = np.where(mask, rng.uniform(8.0, 40.0, disps.shape), disps)
    tg, _ = reproject(poses_gt, disps_gt, intr, ii, jj)
    targets = tg + rng.normal(0, 0.5, tg.shape)
    weights = rng.uniform(0, 1, tg.shape)
    if heuristics:   # dbaf/covisible_graph.py:327-328
        newest = np.maximum(ii, jj) == n - 1
        weights[newest] /= 4.0
        weights /= 10.0
    t1 = n
    kx = np.unique(np.concatenate([np.arange(t0, t1), ii]))
    eta = np.full((len(kx), ht, wd), 0.2 * 1e-6 + 1e-7)
    eta *= rng.uniform(0.5, 2.0, eta.shape)
    sens = np.zeros_like(disps)
    if sensor_frac > 0:
        m = rng.random(disps.shape) < sensor_frac
        sens = np.where(m, disps_gt * np.exp(rng.normal(0, 0.02, disps.shape)), 0.0)
    f32 = np.float32
    return dict(poses=poses.astype(f32), disps=disps.astype(f32), disps_sens=sens.astype(f32),
                intrinsics=np.array(intr, f32), targets=targets.astype(f32), weights=weights.astype(f32),
                eta=eta.astype(f32), ii=ii, jj=jj, t0=int(t0), t1=int(t1), ht=ht, wd=wd,
                poses_gt=poses_gt.astype(f32), disps_gt=disps_gt.astype(f32))


def make_features(config="c2", seed=0, n_frames=None, channels=128, dtype=np.float16):
    """Feature maps N(0,1) [N,C,ht,wd] (dbaf/depth_video.py:64 stores them in half)."""
    ht, wd, P, _, _ = CONFIGS[config]
    rng = np.random.default_rng(4321 + seed)
    n = n_frames or P
    return rng.normal(0, 1, (n, channels, ht, wd)).astype(dtype)


def lookup_coords(window, oob_frac=0.05, seed=0):
    """Lookup coordinates [E,2,ht,wd] (x then y): the GT reprojection, with a fraction of
    pixels thrown outside the image to exercise zero padding."""
    rng = np.random.default_rng(99 + seed)
    c, _ = reproject(window["poses_gt"].astype(np.float64), window["disps_gt"].astype(np.float64),
                     window["intrinsics"].astype(np.float64), window["ii"], window["jj"])
    m = rng.random(c[:, :1].shape) < oob_frac
    c = np.where(m, c + rng.choice([-1.0, 1.0], c.shape) * rng.uniform(40, 200, c.shape), c)
    return c.astype(np.float32)
