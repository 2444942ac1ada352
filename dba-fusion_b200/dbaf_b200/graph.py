"""Frame-graph edge selection on the device: the proximity / NMS sweep of
CovisibleGraph.add_proximity_factors (dbaf/covisible_graph.py:357-441) and the duplicate filter of
CovisibleGraph.__filter_repeated_edges (:61-72), without the reference's per-candidate `.item()` round trips.

    ii, jj = proximity_factors(poses, disps, intrinsics, t, t0, t1, graph_edges, rad=.., nms=.., ...)

returns exactly the edges the reference would pass on to the rest of `add_factors` (after its duplicate
filter).  Everything runs on the current stream; the only host synchronisation is the read of the final count.
"""
from __future__ import annotations

import torch

import droid_backends as db
from droid_backends._lib import c_void_p, check, lib


def _p(t):
    return None if t is None or t.numel() == 0 else c_void_p(t.data_ptr())


def _stream():
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def proximity_candidates(t0, t1, t, skip_edge=(), frontend_window=0, device="cuda"):
    """The candidate pairs of :360-377 as device tensors (no synchronisation: everything follows from the integers).
    Returns ii, jj (int64) and cc, the number of grid entries; the rest are the "skip" pairs."""
    W = t - t1
    k = torch.arange((t - t0) * W, device=device)
    ii, jj = t0 + k // max(W, 1), t1 + k % max(W, 1)
    cc = int(k.numel())
    if len(skip_edge) and cc and (t - 1) - t0 == frontend_window - 1:
        add = [t0 + s for s in skip_edge if t0 + s > 0]
        ii = torch.cat([ii, torch.full((len(add),), t - 1, dtype=torch.int64, device=device)])
        jj = torch.cat([jj, torch.as_tensor(add, dtype=torch.int64, device=device)])
    return ii.contiguous(), jj.contiguous(), cc


def select_edges(d1, d2, cand_ii, cand_jj, cc, ex_ii, ex_jj, t0, t1, t, rad, nms, thresh, stereo, max_factors):
    """dbaf_proximity_select.  Returns (es_ii, es_jj, n_es): fixed-capacity int64 tensors and the device count."""
    dev = cand_ii.device
    n_cand = int(cand_ii.numel())
    cap = lib.dbaf_proximity_base_count(int(t0), int(t), int(rad), int(bool(stereo))) + 2 * cc + 2
    es_ii = torch.empty(cap, dtype=torch.int64, device=dev)
    es_jj = torch.empty(cap, dtype=torch.int64, device=dev)
    n_es = torch.zeros(1, dtype=torch.int32, device=dev)
    work = torch.empty(max(n_cand, 1), dtype=torch.float32, device=dev)
    ex_ii, ex_jj = ex_ii.contiguous(), ex_jj.contiguous()
    check(lib.dbaf_proximity_select(_p(d1), _p(d2), _p(cand_ii), _p(cand_jj), n_cand, cc, _p(ex_ii), _p(ex_jj),
                                    int(ex_ii.numel()), int(t0), int(t1), int(t), int(rad), int(nms), float(thresh),
                                    int(bool(stereo)), int(max_factors), _p(work), _p(es_ii), _p(es_jj), cap, _p(n_es),
                                    _stream()), "dbaf_proximity_select")
    return es_ii, es_jj, n_es


def filter_repeated_edges(ii, jj, n_dev, ex_ii, ex_jj):
    """dbaf_filter_repeated_edges.  Returns (out_ii, out_jj, n_out) like select_edges."""
    dev = ii.device
    n_max = int(ii.numel())
    out_ii, out_jj = torch.empty_like(ii), torch.empty_like(jj)
    n_out = torch.zeros(1, dtype=torch.int32, device=dev)
    ex_ii, ex_jj = ex_ii.contiguous(), ex_jj.contiguous()
    check(lib.dbaf_filter_repeated_edges(_p(ii), _p(jj), _p(n_dev), n_max, _p(ex_ii), _p(ex_jj), int(ex_ii.numel()),
                                         _p(out_ii), _p(out_jj), _p(n_out), _stream()), "dbaf_filter_repeated_edges")
    return out_ii, out_jj, n_out


def proximity_factors(poses, disps, intrinsics, t, t0, t1, active, bad, inactive, rad=2, nms=2, beta=0.25,
                      thresh=16.0, stereo=False, max_factors=48, skip_edge=(), frontend_window=0, filtered=True):
    """add_proximity_factors up to (and, with `filtered`, including) the duplicate filter of add_factors.
    poses [N,7], disps [N,ht,wd], intrinsics [4]: the DepthVideo buffers; t: number of keyframes;
    active / bad / inactive: (ii, jj) pairs of int64 CUDA tensors (self.ii, self.ii_bad, self.ii_inac and their jj).
    Returns (ii, jj): int64 CUDA tensors."""
    dev = poses.device
    ci, cj, cc = proximity_candidates(t0, t1, t, skip_edge, frontend_window, dev)
    p = poses[:t].contiguous()
    d1 = db.frame_distance(p, disps, intrinsics, ci, cj, beta)
    d2 = db.frame_distance(p, disps, intrinsics, cj, ci, beta)
    ex_ii = torch.cat([active[0], bad[0], inactive[0]])
    ex_jj = torch.cat([active[1], bad[1], inactive[1]])
    es_ii, es_jj, n_es = select_edges(d1, d2, ci, cj, cc, ex_ii, ex_jj, t0, t1, t, rad, nms, thresh, stereo, max_factors)
    if filtered:
        es_ii, es_jj, n_es = filter_repeated_edges(es_ii, es_jj, n_es, torch.cat([active[0], inactive[0]]),
                                                   torch.cat([active[1], inactive[1]]))
    n = int(n_es.item())                               # the one host synchronisation of the whole selection
    return es_ii[:n], es_jj[:n]
