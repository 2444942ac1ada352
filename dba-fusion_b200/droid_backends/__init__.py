"""`droid_backends` for B200: same names, argument order and return values as the module
DBA-Fusion builds from src/droid.cpp (PYBIND11_MODULE at src/droid.cpp:297-316), backed by
libdbaf_b200.so through its C ABI.  Put the directory that contains this package on
PYTHONPATH ahead of the reference build and `import droid_backends` resolves here; nothing in
dbaf/ needs to change (callers: dbaf/modules/corr.py:12,19,79,87, dbaf/depth_video.py:255-265,
332,392-397,469-478,527,558, dbaf/dbaf.py:76,79,119,121).

Tensors are torch CUDA tensors; torch only provides memory, the current stream and
(for global BA) the process group.  All kernels are enqueued on the current stream.
Extras beyond the reference surface: corr_volume_pyramid, corr_lookup_pyramid, ba_tables,
ba_info, set_option, set_strict.

Status words.  The kernels never fault on a bad edge index or an `eta` whose row count does not match the
number of depth frames; they raise a flag on the device instead (int32 info[2], info[3]) and skip the work
that depends on it.  Calls that synchronise with the device anyway (BACore.hessian / hessian_gtsam, ba_extend,
ba_host) read the flags and raise RuntimeError like the reference would; the asynchronous `ba` does so only in
strict mode (set_strict(True) or DBAF_STRICT=1), otherwise poll ba_info().
"""
from __future__ import annotations

import ctypes
import os

import torch

from ._lib import BaLayout, c_void_p, check, lib

__all__ = ["set_option", "set_strict", "ba", "ba_extend", "frame_distance", "projmap", "depth_filter", "iproj",
           "altcorr_forward", "altcorr_backward", "corr_index_forward", "corr_index_backward",
           "BACore", "ba2gtsam_matrix", "corr_volume_pyramid", "corr_lookup_pyramid", "corr_volume_status", "ba_tables", "ba_info"]

DBAF_F16, DBAF_F32 = 0, 1
_strict = os.environ.get("DBAF_STRICT", "0") not in ("", "0")


def set_option(key, value):
    """dbaf_set_option: kernel-selection switches for tests and profiling (see include/dbaf_b200.h)."""
    check(lib.dbaf_set_option(str(key).encode(), str(value).encode()), "dbaf_set_option")


def probe_fill(buf, kind):
    """dbaf_probe_fill: overwrite `buf` (a CUDA tensor) with cudaMemsetAsync (kind 0) or 128-bit stores (kind 1)."""
    check(lib.dbaf_probe_fill(_p(buf), buf.numel() * buf.element_size(), int(kind), _stream()), "dbaf_probe_fill")


def banded_solve_launches(n_poses, band):
    """Number of kernels dbaf_ba_solve_banded enqueues for a window of n_poses poses with that block band."""
    return int(lib.dbaf_ba_solve_banded_launches(int(n_poses), int(band)))


def set_strict(on=True):
    """In strict mode `ba` synchronises after the call and raises on a bad edge index or an eta/K mismatch."""
    global _strict
    _strict = bool(on)


_volume_ws = None          # workspace of the most recent indexed / pool volume build: its tail holds the status word


def corr_volume_status():
    """0 if the most recent corr_volume_pyramid_indexed / corr_volume_pyramid_pool call had all its frame indices and
    slots in range, 1 otherwise (the call then wrote nothing).  Synchronises the current stream."""
    if _volume_ws is None:
        return 0
    ws, nbytes = _volume_ws
    st = torch.zeros(1, dtype=torch.int32).pin_memory()
    check(lib.dbaf_corr_volume_status(_p(ws), nbytes, _p(st), _stream()), "dbaf_corr_volume_status")
    torch.cuda.current_stream().synchronize()
    return int(st[0])


def _volume_done(ws, nbytes, what):
    global _volume_ws
    _volume_ws = (ws, nbytes)
    if _strict and corr_volume_status():
        raise RuntimeError(f"{what}: frame index or slot out of range")


def _raise_on_flags(info, what):
    if int(info[2]):
        raise RuntimeError(f"{what}: an edge index in ii/jj lies outside [0, n_frames)")
    if int(info[3]):
        raise RuntimeError(f"{what}: eta has a row count that is neither 1 nor the number of depth frames "
                           f"K = {int(info[0])} (len(unique(cat(arange(t0,t1), ii))))")


def _contig(**named):
    # CHECK_CONTIGUOUS, src/droid.cpp:105-106
    for name, t in named.items():
        if not t.is_contiguous():
            raise RuntimeError(f"{name} must be contiguous")


def _cuda(dtype, **named):
    for name, t in named.items():
        if not t.is_cuda:
            raise RuntimeError(f"{name} must be a CUDA tensor (droid_backends for B200 has no CPU path)")
        if t.dtype != dtype:
            raise RuntimeError(f"{name} must be {dtype}, got {t.dtype}")


def _p(t):
    return None if t is None else c_void_p(t.data_ptr())


def _stream():
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def _layout(N, E, t0, t1, ht, wd):
    L = BaLayout()
    check(lib.dbaf_ba_workspace_layout(N, E, t0, t1, ht, wd, ctypes.byref(L)), "dbaf_ba_workspace_layout")
    return L


def _ba_checks(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1):
    _contig(targets=targets, weights=weights, poses=poses, disps=disps, intrinsics=intrinsics,
            disps_sens=disps_sens, ii=ii, jj=jj)
    _cuda(torch.float32, poses=poses, disps=disps, intrinsics=intrinsics, disps_sens=disps_sens,
          targets=targets, weights=weights, eta=eta)
    _cuda(torch.int64, ii=ii, jj=jj)
    if disps.dim() != 3 or poses.dim() != 2 or poses.shape[1] != 7:
        raise RuntimeError("poses must be [N,7] and disps [N,ht,wd]")
    N, ht, wd = disps.shape
    E = ii.numel()
    if jj.numel() != E or targets.numel() != E * 2 * ht * wd or weights.numel() != targets.numel():
        raise RuntimeError("targets/weights must be [E,2,ht,wd] with E = len(ii) = len(jj)")
    if poses.shape[0] < t1 or disps_sens.shape != disps.shape or not (0 <= t0 <= t1 <= N):
        raise RuntimeError("need 0 <= t0 <= t1 <= N and matching poses/disps/disps_sens")
    return N, ht, wd, E


def _eta_rows(eta, ii, t0, t1, HW):
    eta = eta.contiguous()
    rows = eta.numel() // HW
    if rows * HW != eta.numel() or rows < 1:
        raise RuntimeError("eta must be [K,ht,wd]")
    if rows == 1:
        # a single row broadcasts over all K depth frames; K then has to be counted
        K = int(torch.unique(torch.cat([torch.arange(t0, t1, device=ii.device), ii])).numel())
    else:
        # the reference demands rows == K (broadcast error otherwise, src/droid_kernels.cu:1475); the device
        # verifies it (info[3]).  dz is always allocated with Kcap rows, so a wrong count cannot write out of bounds
        K = rows
    return eta, rows, K


# Workspaces are recycled between calls: `ba` is called many times per keyframe on windows of the same shape.  A
# workspace is only handed out again for the same device, stream and size, and at most two are kept.
_ws_pool = {}


def _take_ws(nbytes, device):
    key = (device.index, torch.cuda.current_stream(device).cuda_stream, int(nbytes))
    lst = _ws_pool.get(key)
    if lst:
        return lst.pop(), key
    return torch.empty(nbytes, dtype=torch.uint8, device=device), key


def _give_ws(key, t):
    if key not in _ws_pool:
        _ws_pool.clear()                   # a new window shape: drop the old one's workspaces
    lst = _ws_pool.setdefault(key, [])
    if len(lst) < 2 and all(x is not t for x in lst):
        lst.append(t)


class _Window:
    """Workspace + dimensions shared by ba / ba_extend / BACore."""

    def __init__(self, poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1,
                 need_eta=True):
        self.N, self.ht, self.wd, self.E = _ba_checks(poses, disps, intrinsics, disps_sens, targets,
                                                      weights, eta, ii, jj, t0, t1)
        self.t0, self.t1, self.P, self.HW = int(t0), int(t1), int(t1 - t0), self.ht * self.wd
        self.args = (poses, disps, intrinsics, disps_sens, targets, weights, ii, jj)
        if need_eta:
            self.eta, self.eta_rows, self.K = _eta_rows(eta, ii, t0, t1, self.HW)
        else:
            self.eta, self.eta_rows, self.K = eta, max(1, eta.numel() // self.HW), 0
        self.L = _layout(self.N, self.E, self.t0, self.t1, self.ht, self.wd)
        self.ws, self._ws_key = _take_ws(self.L.total_bytes, poses.device)
        self.n = 6 * self.P
        self._pin = None

    def release(self):
        """Return the workspace to the pool (the tensor stays referenced here, so ba_info() keeps working until
        the next call of the same shape overwrites it)."""
        if self._ws_key is not None:
            _give_ws(self._ws_key, self.ws)
            self._ws_key = None

    # views into the workspace ------------------------------------------------
    def _view(self, off, count, dtype):
        nbytes = count * torch.empty((), dtype=dtype).element_size()
        return self.ws[off:off + nbytes].view(dtype)

    def H(self):
        return self._view(self.L.H, self.n * self.n, torch.float64).view(self.n, self.n)

    def v(self):
        return self._view(self.L.v, self.n, torch.float64)

    def Hv(self):
        """H and v as one contiguous float64 span (what global BA all-reduces)."""
        return self._view(self.L.H, self.n * self.n + self.n, torch.float64)

    def dx(self):
        return self._view(self.L.dx, self.n, torch.float32).view(self.P, 6)

    def info(self):
        return self._view(self.L.info, 16, torch.int32)

    def export_host(self, src=None):
        """H|v (or `src`, a float64 device tensor of the same size) and the status words -> pinned host memory
        with one stream synchronisation.  Returns the float64 host tensor; info lands in self._info_host."""
        n2 = self.n * self.n + self.n
        if getattr(self, "_pin", None) is None:
            self._pin = torch.empty(n2, dtype=torch.float64).pin_memory()
            self._info_host = torch.zeros(16, dtype=torch.int32).pin_memory()
        self._pin.copy_(self.Hv() if src is None else src, non_blocking=True)
        self._info_host.copy_(self.info(), non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return self._pin

    # stages ------------------------------------------------------------------
    def prepare(self):
        poses, disps, intr, sens, tg, wt, ii, jj = self.args
        check(lib.dbaf_ba_prepare(_p(self.ws), _p(ii), _p(jj), self.E, self.N, self.t0, self.t1,
                                  self.ht, self.wd, _stream()), "dbaf_ba_prepare")

    def linearize(self, alpha, motion_only=False, zero_first=True, k_lo=0, k_hi=-1):
        poses, disps, intr, sens, tg, wt, ii, jj = self.args
        check(lib.dbaf_ba_linearize(_p(self.ws), _p(poses), _p(disps), _p(intr), _p(sens), _p(tg), _p(wt),
                                    _p(self.eta), self.eta_rows, _p(ii), _p(jj), self.E, self.N, self.t0,
                                    self.t1, self.ht, self.wd, float(alpha), int(motion_only),
                                    int(zero_first), int(k_lo), int(k_hi), _stream()), "dbaf_ba_linearize")

    def zero_system(self, band=0):
        """H, v = 0 ahead of linearize(zero_first=False); with a band hint only the band of H is cleared."""
        check(lib.dbaf_ba_zero_system(_p(self.ws), self.t0, self.t1, int(band), _stream()), "dbaf_ba_zero_system")

    def band_strip(self, band, strip=None, pack=True):
        """The band of H, and v, as one dense float64 strip (pack) or back from it (pack=False)."""
        if strip is None:
            strip = torch.empty(lib.dbaf_ba_band_strip_doubles(self.P, int(band)), dtype=torch.float64, device=self.ws.device)
        check(lib.dbaf_ba_band_strip(_p(self.ws), self.t0, self.t1, int(band), _p(strip), int(pack), _stream()),
              "dbaf_ba_band_strip")
        return strip

    def solve(self, lm, ep, prior=None, band=0, prior_shape=None):
        if band and prior is None:
            check(lib.dbaf_ba_solve_banded(_p(self.ws), self.t0, self.t1, float(lm), float(ep), int(band),
                                           _stream()), "dbaf_ba_solve_banded")
            return
        pr, pc = (0, 0) if prior is None else (prior_shape if prior_shape is not None else prior.shape)
        check(lib.dbaf_ba_solve(_p(self.ws), self.t0, self.t1, float(lm), float(ep), _p(prior),
                                int(pr), int(pc), _stream()), "dbaf_ba_solve")

    def retract(self, dx_in=None, update_poses=True, update_disps=True, k_lo=0, k_hi=-1,
                want_dx=True, want_dz=True):
        poses, disps, intr, sens, tg, wt, ii, jj = self.args
        dx_out = torch.empty(self.P, 6, dtype=torch.float32, device=poses.device) if want_dx else None
        dz_out = (torch.empty(max(self.K, int(self.L.Kcap)), self.HW, dtype=torch.float32, device=poses.device)
                  if (want_dz and update_disps) else None)
        check(lib.dbaf_ba_retract(_p(self.ws), _p(poses), _p(disps), _p(dx_in), _p(ii), _p(jj), self.E,
                                  self.N, self.t0, self.t1, self.ht, self.wd, int(update_poses),
                                  int(update_disps), int(k_lo), int(k_hi), _p(dx_out), _p(dz_out),
                                  _stream()), "dbaf_ba_retract")
        return dx_out, (dz_out[:self.K] if dz_out is not None else None)

    def schur(self, k_lo=0, k_hi=-1):
        """dbaf_ba_schur: the Schur stage alone, on the E, Q, w the workspace holds."""
        jj = self.args[7]
        check(lib.dbaf_ba_schur(_p(self.ws), _p(jj), self.E, self.N, self.t0, self.t1, self.ht, self.wd,
                                int(k_lo), int(k_hi), _stream()), "dbaf_ba_schur")

    def check_flags(self, what, synced=False):
        _raise_on_flags(self._info_host if synced else self.info().cpu(), what)


_last_window = None


def ba_info():
    """int32[16] status words of the most recent ba / ba_extend / BACore window (forces a
    device synchronisation): [0]=K, [1]=#failed factorisations, [2]=bad index, [3]=eta row
    mismatch, [4]=#Schur rows."""
    if _last_window is None:
        return None
    return _last_window.info().cpu()


def ba(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj,
       t0, t1, iterations, lm, ep, motion_only):
    """droid_backends.ba (src/droid.cpp:109-138 -> ba_cuda, src/droid_kernels.cu:1394-1512).
    Updates poses[t0:t1] and disps[kx] in place; returns [dx, dz] of the last iteration
    (dz is None for motion_only, both None when iterations == 0)."""
    global _last_window
    w = _Window(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1,
                need_eta=not motion_only)
    _last_window = w
    if iterations <= 0:
        return [None, None]
    dx = torch.empty(w.P, 6, dtype=torch.float32, device=poses.device)
    dz = None if motion_only else torch.empty(max(w.K, int(w.L.Kcap)), w.HW, dtype=torch.float32,
                                              device=poses.device)
    poses_, disps_, intr, sens, tg, wt, ii_, jj_ = w.args
    check(lib.dbaf_ba(_p(w.ws), _p(poses_), _p(disps_), _p(intr), _p(sens), _p(tg), _p(wt), _p(w.eta),
                      w.eta_rows, _p(ii_), _p(jj_), w.E, w.N, w.t0, w.t1, w.ht, w.wd, int(iterations),
                      float(lm), float(ep), int(bool(motion_only)), _p(dx), _p(dz), _stream()), "dbaf_ba")
    if _strict:
        w.check_flags("ba")
    w.release()
    return [dx, None if dz is None else dz[:w.K]]


_host_cache = {}


def ba_host(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj,
            t0, t1, iterations, lm, ep, motion_only, device=None):
    """`ba` for a window that lives in HOST memory (C entry point dbaf_ba_host; no reference counterpart, the
    reference only takes CUDA tensors).  All arguments are CPU tensors with the shapes and dtypes of `ba`, ideally
    pinned; poses[t0:t1] and disps are updated in place, dx [P,6] (CPU, pinned) is returned.  The inputs are
    copied to a cached device staging buffer, the iterations run there, the results are copied back, and the call
    returns after synchronising the stream."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    for name, t, dt in (("poses", poses, torch.float32), ("disps", disps, torch.float32),
                        ("intrinsics", intrinsics, torch.float32), ("disps_sens", disps_sens, torch.float32),
                        ("targets", targets, torch.float32), ("weights", weights, torch.float32),
                        ("eta", eta, torch.float32), ("ii", ii, torch.int64), ("jj", jj, torch.int64)):
        if t.is_cuda or t.dtype != dt or not t.is_contiguous():
            raise RuntimeError(f"{name} must be a contiguous CPU tensor of {dt}")
    if disps.dim() != 3 or poses.dim() != 2 or poses.shape[1] != 7:
        raise RuntimeError("poses must be [N,7] and disps [N,ht,wd]")
    N, ht, wd = disps.shape
    E, HW = ii.numel(), ht * wd
    if jj.numel() != E or targets.numel() != E * 2 * HW or weights.numel() != targets.numel():
        raise RuntimeError("targets/weights must be [E,2,ht,wd] with E = len(ii) = len(jj)")
    if poses.shape[0] < t1 or disps_sens.shape != disps.shape or not (0 <= t0 <= t1 <= N):
        raise RuntimeError("need 0 <= t0 <= t1 <= N and matching poses/disps/disps_sens")
    rows = eta.numel() // HW
    if rows * HW != eta.numel() or (rows < 1 and not motion_only):
        raise RuntimeError("eta must be [K,ht,wd]")
    key = (dev.index, N, E, int(t0), int(t1), ht, wd, rows)
    ent = _host_cache.get(key)
    if ent is None:
        L = _layout(N, E, int(t0), int(t1), ht, wd)
        nstage = lib.dbaf_ba_host_staging_bytes(N, E, int(t0), int(t1), ht, wd, rows)
        ent = (torch.empty(L.total_bytes, dtype=torch.uint8, device=dev),
               torch.empty(nstage, dtype=torch.uint8, device=dev), nstage,
               torch.empty(max(int(t1 - t0), 1), 6, dtype=torch.float32).pin_memory(),
               torch.zeros(16, dtype=torch.int32).pin_memory())
        _host_cache.clear()            # one window shape at a time is the common case; do not hoard device memory
        _host_cache[key] = ent
    ws, stage, nstage, dx, info = ent
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream()
        check(lib.dbaf_ba_host(_p(ws), _p(stage), nstage, _p(poses), _p(disps), _p(intrinsics), _p(disps_sens),
                               _p(targets), _p(weights), _p(eta), rows, _p(ii), _p(jj), E, N, int(t0), int(t1), ht,
                               wd, int(iterations), float(lm), float(ep), int(bool(motion_only)), _p(dx),
                               _p(info), c_void_p(stream.cuda_stream)), "dbaf_ba_host")
        stream.synchronize()
    if iterations > 0:
        _raise_on_flags(info, "ba_host")
    return dx[:int(t1 - t0)]


def _export(w, H, v):
    """Copy the undamped reduced system to the caller's float64 CPU tensors
    (src/droid_kernels.cu:1632-1640, 1889-1897: only H.size(0) x H.size(1) entries are written)."""
    if H.dtype != torch.float64 or v.dtype != torch.float64 or H.is_cuda or v.is_cuda:
        raise RuntimeError("H and v must be float64 CPU tensors")
    r, c = H.shape
    if r > w.n or c > w.n:
        raise RuntimeError(f"H is larger than the {w.n}x{w.n} reduced camera system")
    hv = w.export_host()                   # one D2H copy into pinned memory, synchronises the stream
    w.check_flags("hessian export", synced=True)
    H.copy_(hv[:w.n * w.n].view(w.n, w.n)[:r, :c])
    v[:r].copy_(hv[w.n * w.n:][:r])


def ba_extend(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, H, v, A_prior,
              t0, t1, iterations, lm, ep, motion_only, skip_solve):
    """droid_backends.ba_extend (src/droid.cpp:140-178 -> ba_cuda_extend,
    src/droid_kernels.cu:1514-1678)."""
    global _last_window
    _contig(H=H, v=v, A_prior=A_prior)
    w = _Window(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1,
                need_eta=not motion_only)
    _last_window = w
    dx = dz = None
    if iterations <= 0:
        return [dx, dz]
    w.prepare()
    prior = None
    for _ in range(int(iterations)):
        w.linearize(0.05, motion_only=motion_only)
        if motion_only:
            w.solve(lm, ep)
            dx, _ = w.retract(update_disps=False)
            continue
        _export(w, H, v)
        if skip_solve:
            return [dx, dz]
        if prior is None:
            if A_prior.dtype != torch.float64 or A_prior.dim() != 2:
                raise RuntimeError("A_prior must be a 2-D float64 tensor")
            prior_shape = tuple(A_prior.shape)
            # a non-null pointer even for an empty prior: the reference then still solves with Ad + Adprior = 2 Ad
            prior = (A_prior.to(poses.device).contiguous() if A_prior.numel() > 0
                     else torch.zeros(1, dtype=torch.float64, device=poses.device))
        w.solve(lm, ep, prior, prior_shape=prior_shape)
        dx, dz = w.retract()
    return [dx, dz]


class BACore:
    """Stateful BA object (src/bacore.h:4-70, src/droid_kernels.cu:1786-1956): `hessian` exports
    the undamped reduced camera system for the GTSAM factor graph, `retract` applies the
    externally solved pose update plus the depth back-substitution."""

    def __init__(self):
        self._w = None

    def init(self, poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj,
             t0, t1, iterations, lm, ep, motion_only):
        # `iterations` and `motion_only` are accepted and ignored, as in the reference (:1802-1814)
        global _last_window
        self._w = _Window(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1)
        _last_window = self._w
        self.lm, self.ep = float(lm), float(ep)
        self.dx = None
        self.dz = None
        self._w.prepare()

    def hessian(self, H, v):
        w = self._w
        w.linearize(0.001)                 # alpha = 0.001, src/droid_kernels.cu:1872
        _export(w, H, v)

    def hessian_gtsam(self, H, v, A):
        """`hessian` followed by gtsam.BA2GTSAM (dbaf/depth_video.py:20-29, call site :527-530) in one go:
        H <- J^T H J, v <- J^T v with J = blockdiag(A), computed on the device before the copy to the caller's
        float64 CPU tensors.  A: [6,6] float64 (the row-swapped -Ad(Tbc^-1) of the reference)."""
        w = self._w
        if H.dtype != torch.float64 or v.dtype != torch.float64 or H.is_cuda or v.is_cuda:
            raise RuntimeError("H and v must be float64 CPU tensors")
        if tuple(H.shape) != (w.n, w.n) or v.numel() != w.n:
            raise RuntimeError("hessian_gtsam expects the full reduced system")
        w.linearize(0.001)                 # alpha = 0.001, src/droid_kernels.cu:1872
        dev = w.ws.device
        Ad = torch.as_tensor(A, dtype=torch.float64).reshape(6, 6).contiguous().to(dev)
        out = torch.empty(w.n * w.n + w.n, dtype=torch.float64, device=dev)
        check(lib.dbaf_ba_congruence(_p(w.ws), w.t0, w.t1, _p(Ad), _p(out), c_void_p(out.data_ptr() + 8 * w.n * w.n),
                                     _stream()), "dbaf_ba_congruence")
        host = w.export_host(out)          # one D2H copy into pinned memory, synchronises the stream
        w.check_flags("hessian_gtsam", synced=True)
        H.copy_(host[:w.n * w.n].view(w.n, w.n))
        v.reshape(-1).copy_(host[w.n * w.n:])

    def optimize(self, H, v):
        w = self._w
        if H.shape != (w.n, w.n) or v.numel() != w.n:
            raise RuntimeError("optimize expects the full reduced system")
        hv = torch.cat([H.reshape(-1).to(torch.float64), v.reshape(-1).to(torch.float64)])
        w.Hv().copy_(hv.to(w.ws.device))
        w.solve(self.lm, self.ep)
        self.dx = w.dx().clone()

    def retract(self, dx):
        w = self._w
        if dx.is_cuda or dx.dtype != torch.float64:
            raise RuntimeError("dx must be a float64 CPU tensor")
        dxd = dx.reshape(-1)[:w.n].to(torch.float32).to(w.ws.device).contiguous()
        self.dx, self.dz = w.retract(dx_in=dxd)
        return [self.dx, self.dz]


def ba2gtsam_matrix(Tbc):
    """The 6x6 matrix A of BA2GTSAM (dbaf/depth_video.py:21-23) from the body<-camera extrinsic Tbc, given as a 4x4
    homogeneous matrix (what gtsam.Pose3.matrix() returns): A0 = -Adjoint(Tbc^-1) in GTSAM's (rotation, translation)
    tangent order, i.e. for T = (R, t) Adjoint(T) = [[R, 0], [skew(t) R, R]], then the two row triples swapped.
    Feed the result to BACore.hessian_gtsam."""
    import numpy as np
    T = np.asarray(Tbc, np.float64).reshape(4, 4)
    R, t = T[:3, :3], T[:3, 3]
    Ri, ti = R.T, -R.T @ t                                  # Tbc^-1
    sk = np.array([[0.0, -ti[2], ti[1]], [ti[2], 0.0, -ti[0]], [-ti[1], ti[0], 0.0]])
    A0 = np.zeros((6, 6))
    A0[:3, :3] = Ri
    A0[3:, :3] = sk @ Ri
    A0[3:, 3:] = Ri
    A0 = -A0
    return np.concatenate([A0[3:6], A0[0:3]], axis=0)


def ba_tables(ii, jj, t0, t1, n_frames, ht=1, wd=1):
    """Graph bookkeeping tables built on the device (kx, kk_exp, csr_ptr/csr_idx, srow_ptr/srow_idx)
    as int64 CPU tensors -- the tables ba_cuda derives at src/droid_kernels.cu:1416-1424,
    993-1031 and 1316-1328."""
    _cuda(torch.int64, ii=ii, jj=jj)
    E = ii.numel()
    L = _layout(n_frames, E, t0, t1, ht, wd)
    ws = torch.empty(L.total_bytes, dtype=torch.uint8, device=ii.device)
    check(lib.dbaf_ba_prepare(_p(ws), _p(ii.contiguous()), _p(jj.contiguous()), E, n_frames, t0, t1, ht, wd,
                              _stream()), "dbaf_ba_prepare")
    P = t1 - t0

    def iv(off, n):
        return ws[off:off + 4 * n].view(torch.int32).cpu().to(torch.int64)

    info = iv(L.info, 16)
    K = int(info[0])
    ptr = iv(L.csr_ptr, K + 1)
    sp = iv(L.srow_ptr, K + 1)
    return dict(info=info, kx=iv(L.kx, K), kk_exp=iv(L.kk_exp, P + E), csr_ptr=ptr,
                csr_idx=iv(L.csr_idx, int(ptr[-1])), srow_ptr=sp, srow_idx=iv(L.srow_idx, int(sp[-1])))


# ----------------------------------------------------------------------- geometry --
def frame_distance(poses, disps, intrinsics, ii, jj, beta):
    """src/droid.cpp:181-197."""
    _contig(poses=poses, disps=disps, intrinsics=intrinsics, ii=ii, jj=jj)
    _cuda(torch.float32, poses=poses, disps=disps, intrinsics=intrinsics)
    _cuda(torch.int64, ii=ii, jj=jj)
    M = ii.numel()
    dist = torch.empty(M, dtype=torch.float32, device=poses.device)
    check(lib.dbaf_frame_distance(_p(poses), _p(disps), _p(intrinsics), _p(ii), _p(jj), M, disps.shape[1],
                                  disps.shape[2], float(beta), _p(dist), _stream()), "dbaf_frame_distance")
    return dist


def projmap(poses, disps, intrinsics, ii, jj):
    """src/droid.cpp:200-215.  Returns [coords [E,ht,wd,3], valid [E,ht,wd,1]]."""
    _contig(poses=poses, disps=disps, intrinsics=intrinsics, ii=ii, jj=jj)
    _cuda(torch.float32, poses=poses, disps=disps, intrinsics=intrinsics)
    _cuda(torch.int64, ii=ii, jj=jj)
    E, ht, wd = ii.numel(), disps.shape[1], disps.shape[2]
    coords = torch.empty(E, ht, wd, 3, dtype=torch.float32, device=poses.device)
    valid = torch.empty(E, ht, wd, 1, dtype=torch.float32, device=poses.device)
    check(lib.dbaf_projmap(_p(poses), _p(disps), _p(intrinsics), _p(ii), _p(jj), E, ht, wd, _p(coords),
                           _p(valid), _stream()), "dbaf_projmap")
    return [coords, valid]


def iproj(poses, disps, intrinsics):
    """src/droid.cpp:218-227.  Returns points [N,ht,wd,3] for N = disps.shape[0]."""
    _contig(poses=poses, disps=disps, intrinsics=intrinsics)
    _cuda(torch.float32, poses=poses, disps=disps, intrinsics=intrinsics)
    N, ht, wd = disps.shape
    pts = torch.empty(N, ht, wd, 3, dtype=torch.float32, device=disps.device)
    check(lib.dbaf_iproj(_p(poses), _p(disps), _p(intrinsics), N, ht, wd, _p(pts), _stream()), "dbaf_iproj")
    return pts


def depth_filter(poses, disps, intrinsics, ix, thresh):
    """src/droid.cpp:281-295.  Returns counter [M,ht,wd]."""
    _contig(poses=poses, disps=disps, intrinsics=intrinsics, ix=ix, thresh=thresh)
    _cuda(torch.float32, poses=poses, disps=disps, intrinsics=intrinsics, thresh=thresh)
    _cuda(torch.int64, ix=ix)
    N, ht, wd = disps.shape
    M = ix.numel()
    counter = torch.empty(M, ht, wd, dtype=torch.float32, device=disps.device)
    check(lib.dbaf_depth_filter(_p(poses), _p(disps), _p(intrinsics), _p(ix), _p(thresh), M, N, ht, wd,
                                _p(counter), _stream()), "dbaf_depth_filter")
    return counter


# -------------------------------------------------------------------- correlation --
def _corr_dtype(t, name):
    if t.dtype == torch.float16:
        return DBAF_F16
    if t.dtype == torch.float32:
        return DBAF_F32
    raise RuntimeError(f"{name}: only float16 and float32 are supported, got {t.dtype}")


def corr_index_forward(volume, coords, radius):
    """src/droid.cpp:231-239.  volume [n,h1,w1,h2,w2], coords [n,2,h1,w1] -> [corr [n,2r+1,2r+1,h1,w1]]."""
    _contig(volume=volume, coords=coords)
    _cuda(torch.float32, coords=coords)
    if not volume.is_cuda:
        raise RuntimeError("volume must be a CUDA tensor")
    dt = _corr_dtype(volume, "volume")
    n, h1, w1, h2, w2 = volume.shape
    rd = 2 * radius + 1
    corr = torch.empty(n, rd, rd, h1, w1, dtype=volume.dtype, device=volume.device)
    check(lib.dbaf_corr_index_forward(_p(volume), dt, _p(coords), n, h1, w1, h2, w2, int(radius), _p(corr),
                                      _stream()), "dbaf_corr_index_forward")
    return [corr]


def corr_index_backward(volume, coords, corr_grad, radius):
    """src/droid.cpp:241-252.  Returns [volume_grad]."""
    _contig(volume=volume, coords=coords, corr_grad=corr_grad)
    _cuda(torch.float32, coords=coords)
    dt = _corr_dtype(volume, "volume")
    if corr_grad.dtype != volume.dtype:
        raise RuntimeError("corr_grad must have the dtype of volume")
    n, h1, w1, h2, w2 = volume.shape
    grad = torch.empty_like(volume)
    check(lib.dbaf_corr_index_backward(_p(coords), _p(corr_grad), dt, n, h1, w1, h2, w2, int(radius),
                                       _p(grad), _stream()), "dbaf_corr_index_backward")
    return [grad]


def altcorr_forward(fmap1, fmap2, coords, radius):
    """src/droid.cpp:254-264.  fmap1 [B,H1,W1,C], fmap2 [B,H2,W2,C], coords [B,S,H1,W1,2]."""
    _contig(fmap1=fmap1, fmap2=fmap2, coords=coords)
    if fmap1.dtype != torch.float32:
        # the reference dispatches on dtype but DBA-Fusion only ever passes float32 (corr.py:120)
        raise RuntimeError("altcorr_forward: float32 feature maps only")
    _cuda(torch.float32, fmap1=fmap1, fmap2=fmap2, coords=coords)
    B, S, H1, W1, _ = coords.shape
    _, H2, W2, C = fmap2.shape
    rd = 2 * radius + 1
    corr = torch.empty(B, S, rd * rd, H1, W1, dtype=torch.float32, device=fmap1.device)
    check(lib.dbaf_altcorr_forward(_p(fmap1), _p(fmap2), _p(coords), B, S, H1, W1, H2, W2, C, int(radius),
                                   _p(corr), _stream()), "dbaf_altcorr_forward")
    return [corr]


def altcorr_backward(fmap1, fmap2, coords, corr_grad, radius):
    """src/droid.cpp:266-279.  Returns [fmap1_grad, fmap2_grad, coords_grad (zeros)]."""
    _contig(fmap1=fmap1, fmap2=fmap2, coords=coords, corr_grad=corr_grad)
    _cuda(torch.float32, fmap1=fmap1, fmap2=fmap2, coords=coords, corr_grad=corr_grad)
    B, S, H1, W1, _ = coords.shape
    _, H2, W2, C = fmap2.shape
    g1 = torch.empty_like(fmap1)
    g2 = torch.empty_like(fmap2)
    gc = torch.empty_like(coords)
    check(lib.dbaf_altcorr_backward(_p(fmap1), _p(fmap2), _p(coords), _p(corr_grad), B, S, H1, W1, H2, W2, C,
                                    int(radius), _p(g1), _p(g2), _p(gc), _stream()), "dbaf_altcorr_backward")
    return [g1, g2, gc]


def reproject(poses, disps, intrinsics, ii, jj):
    """DepthVideo.reproject (dbaf/depth_video.py:221-229 -> geom/projective_ops.py:96-125, no Jacobians) in one
    kernel.  poses [N,7], disps [N,ht,wd], intrinsics [4] or [N,4], ii/jj [E] -> coords [1,E,ht,wd,2] and
    valid [1,E,ht,wd,1], the shapes the reference returns for its batch of one."""
    _contig(poses=poses, disps=disps, intrinsics=intrinsics, ii=ii, jj=jj)
    _cuda(torch.float32, poses=poses, disps=disps, intrinsics=intrinsics)
    _cuda(torch.int64, ii=ii, jj=jj)
    N, ht, wd = disps.shape
    per_frame = intrinsics.dim() == 2
    if intrinsics.numel() != (4 * N if per_frame else 4) or ii.shape != jj.shape:
        raise RuntimeError("intrinsics must be [4] or [N,4]; ii and jj must have the same length")
    E = ii.numel()
    coords = torch.empty(1, E, ht, wd, 2, dtype=torch.float32, device=poses.device)
    valid = torch.empty(1, E, ht, wd, 1, dtype=torch.float32, device=poses.device)
    check(lib.dbaf_reproject(_p(poses), _p(disps), _p(intrinsics), int(per_frame), _p(ii), _p(jj), E, ht, wd,
                             _p(coords), _p(valid), _stream()), "dbaf_reproject")
    return coords, valid


def corr_volume_pyramid(fmap1, fmap2, num_levels=4):
    """Native CorrBlock.__init__ (dbaf/modules/corr.py:24-38,63-71): fmap1, fmap2 [n,C,ht,wd]
    float16 -> list of `num_levels` tensors [n,ht,wd,ht>>l,wd>>l] float16."""
    _contig(fmap1=fmap1, fmap2=fmap2)
    _cuda(torch.float16, fmap1=fmap1, fmap2=fmap2)
    n, C, ht, wd = fmap1.shape
    if fmap2.shape != fmap1.shape:
        raise RuntimeError("fmap1 and fmap2 must have the same shape")
    outs = [torch.empty(n, ht, wd, ht >> l, wd >> l, dtype=torch.float16, device=fmap1.device)
            for l in range(num_levels)]
    nbytes = lib.dbaf_corr_volume_workspace_bytes(n, C, ht, wd)
    ws = torch.empty(max(1, nbytes), dtype=torch.uint8, device=fmap1.device)
    arr = (c_void_p * num_levels)(*[o.data_ptr() for o in outs])
    check(lib.dbaf_corr_volume_pyramid(_p(fmap1), _p(fmap2), n, C, ht, wd, num_levels, arr, _p(ws), nbytes,
                                       _stream()), "dbaf_corr_volume_pyramid")
    return outs


def corr_volume_pyramid_indexed(fmaps1, fmaps2, ii, jj, num_levels=4):
    """CorrBlock(video.fmaps[ii,0], video.fmaps[jj,c]) (dbaf/covisible_graph.py:129-131) without the gather:
    fmaps1 [n1,C,ht,wd], fmaps2 [n2,C,ht,wd] float16 frame stacks (they may be the same tensor), ii/jj edge
    lists -> list of `num_levels` tensors [E,ht,wd,ht>>l,wd>>l] float16, bit-identical to
    corr_volume_pyramid(fmaps1[ii], fmaps2[jj]).  An index outside its stack makes the call write nothing; it is
    reported by corr_volume_status() (or raised here in strict mode) -- the call itself never synchronises."""
    _contig(fmaps1=fmaps1, fmaps2=fmaps2)
    _cuda(torch.float16, fmaps1=fmaps1, fmaps2=fmaps2)
    n1, C, ht, wd = fmaps1.shape
    n2 = fmaps2.shape[0]
    if fmaps2.shape[1:] != fmaps1.shape[1:]:
        raise RuntimeError("fmaps1 and fmaps2 must agree in channels and size")
    if ii.shape != jj.shape or ii.dim() != 1:
        raise RuntimeError("ii and jj must be 1-D and of equal length")
    ii32 = ii.to(device=fmaps1.device, dtype=torch.int32).contiguous()
    jj32 = jj.to(device=fmaps1.device, dtype=torch.int32).contiguous()
    E = ii32.numel()
    outs = [torch.empty(E, ht, wd, ht >> l, wd >> l, dtype=torch.float16, device=fmaps1.device)
            for l in range(num_levels)]
    nbytes = lib.dbaf_corr_volume_indexed_workspace_bytes(n1, n2, C, ht, wd)
    ws = torch.empty(max(1, nbytes), dtype=torch.uint8, device=fmaps1.device)
    arr = (c_void_p * num_levels)(*[o.data_ptr() for o in outs])
    check(lib.dbaf_corr_volume_pyramid_indexed(_p(fmaps1), n1, _p(fmaps2), n2, _p(ii32), _p(jj32), E, C, ht, wd,
                                               num_levels, arr, _p(ws), nbytes, _stream()),
          "dbaf_corr_volume_pyramid_indexed")
    _volume_done(ws, nbytes, "corr_volume_pyramid_indexed")
    return outs


def corr_volume_pyramid_pool(fmaps1, fmaps2, ii, jj, slots, pool):
    """Build the pyramids of the edges (ii[e], jj[e]) straight into slots `slots[e]` of `pool` (a list of
    [capacity, ht, wd, ht>>l, wd>>l] float16 tensors): C entry point dbaf_corr_volume_pyramid_pool, see
    dbaf_b200.corr.CorrPool."""
    _contig(fmaps1=fmaps1, fmaps2=fmaps2, **{f"pool[{l}]": t for l, t in enumerate(pool)})
    _cuda(torch.float16, fmaps1=fmaps1, fmaps2=fmaps2, **{f"pool[{l}]": t for l, t in enumerate(pool)})
    n1, C, ht, wd = fmaps1.shape
    n2 = fmaps2.shape[0]
    cap = pool[0].shape[0]
    for l, t in enumerate(pool):
        if tuple(t.shape) != (cap, ht, wd, ht >> l, wd >> l):
            raise RuntimeError(f"pool level {l} has the wrong shape")
    dev = fmaps1.device
    ii32 = ii.to(device=dev, dtype=torch.int32).contiguous()
    jj32 = jj.to(device=dev, dtype=torch.int32).contiguous()
    sl32 = slots.to(device=dev, dtype=torch.int32).contiguous()
    E = ii32.numel()
    if jj32.numel() != E or sl32.numel() != E:
        raise RuntimeError("ii, jj and slots must have the same length")
    nbytes = lib.dbaf_corr_volume_indexed_workspace_bytes(n1, n2, C, ht, wd)
    ws = torch.empty(max(1, nbytes), dtype=torch.uint8, device=dev)
    arr = (c_void_p * len(pool))(*[t.data_ptr() for t in pool])
    check(lib.dbaf_corr_volume_pyramid_pool(_p(fmaps1), n1, _p(fmaps2), n2, _p(ii32), _p(jj32), _p(sl32), E, cap, C, ht,
                                            wd, len(pool), arr, _p(ws), nbytes, _stream()),
          "dbaf_corr_volume_pyramid_pool")
    _volume_done(ws, nbytes, "corr_volume_pyramid_pool")


def corr_lookup_pyramid_pool(pool, coords, slots, radius=3):
    """corr_lookup_pyramid over a slot pool: edge e (coords [n,2,h1,w1]) reads volume slots[e] of every level."""
    _contig(coords=coords, slots=slots, **{f"pool[{i}]": t for i, t in enumerate(pool)})
    _cuda(torch.float32, coords=coords)
    _cuda(torch.int32, slots=slots)
    dt = _corr_dtype(pool[0], "pool")
    cap, h1, w1, h2, w2 = pool[0].shape
    n = coords.shape[0]
    if slots.numel() != n or tuple(coords.shape) != (n, 2, h1, w1):
        raise RuntimeError("coords must be [n,2,h1,w1] with one slot per edge")
    rd = 2 * radius + 1
    out = torch.empty(n, len(pool) * rd * rd, h1, w1, dtype=pool[0].dtype, device=coords.device)
    arr = (c_void_p * len(pool))(*[t.data_ptr() for t in pool])
    check(lib.dbaf_corr_lookup_pyramid_pool(arr, len(pool), dt, _p(coords), _p(slots), n, cap, h1, w1, h2, w2,
                                            int(radius), _p(out), _stream()), "dbaf_corr_lookup_pyramid_pool")
    return out


def corr_lookup_pyramid(pyramid, coords, radius=3):
    """All levels of CorrBlock.__call__ (dbaf/modules/corr.py:40-50) in one launch.
    pyramid[l] [n,h1,w1,h2>>l,w2>>l]; coords [n,2,h1,w1] (level-0 pixels) ->
    [n, levels*(2r+1)^2, h1, w1], i.e. torch.cat(out_pyramid, dim=2) without the batch dim."""
    _contig(coords=coords, **{f"pyramid[{i}]": t for i, t in enumerate(pyramid)})
    _cuda(torch.float32, coords=coords)
    dt = _corr_dtype(pyramid[0], "pyramid")
    n, h1, w1, h2, w2 = pyramid[0].shape
    for l, t in enumerate(pyramid):
        if tuple(t.shape) != (n, h1, w1, h2 >> l, w2 >> l) or t.dtype != pyramid[0].dtype:
            raise RuntimeError(f"pyramid level {l} has the wrong shape or dtype")
    rd = 2 * radius + 1
    out = torch.empty(n, len(pyramid) * rd * rd, h1, w1, dtype=pyramid[0].dtype, device=coords.device)
    arr = (c_void_p * len(pyramid))(*[t.data_ptr() for t in pyramid])
    check(lib.dbaf_corr_lookup_pyramid(arr, len(pyramid), dt, _p(coords), n, h1, w1, h2, w2, int(radius),
                                       _p(out), _stream()), "dbaf_corr_lookup_pyramid")
    return out
