"""Frame-graph edge selection: CovisibleGraph.add_proximity_factors and
__filter_repeated_edges (dbaf/covisible_graph.py:357-441, 61-72).

The golden vectors in tests/golden/ref_proximity.npz were produced by the reference's own Python methods
(tests/golden/make_golden_proximity.py).  CPU: the numpy restatement reproduces them.  GPU: the kernels reproduce
them and the restatement on further random windows -- edge lists bit for bit, order included."""
import os

import numpy as np
import pytest
import torch

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_proximity.npz")


def _cases():
    g = np.load(GOLD)
    for c in range(int(g["n_cases"])):
        p = f"case{c}_"
        t, t0, t1, rad, nms, stereo, mf, win = g[p + "params"].tolist()
        yield dict(t=t, t0=t0, t1=t1, rad=rad, nms=nms, stereo=bool(stereo), max_factors=mf, window=win,
                   thresh=float(g[p + "thresh"]), skip=g[p + "skip"].tolist(), ex=g[p + "ex"], act=g[p + "act"],
                   d=g[p + "d"], cand_ii=g[p + "cand_ii"], cand_jj=g[p + "cand_jj"], es_ii=g[p + "es_ii"],
                   es_jj=g[p + "es_jj"], f_ii=g[p + "f_ii"], f_jj=g[p + "f_jj"])


def test_oracle_reproduces_the_reference_edge_lists(oracle):
    n = 0
    for c in _cases():
        ii, jj, cc = oracle.proximity_candidates(c["t0"], c["t1"], c["t"], c["skip"], c["window"])
        assert np.array_equal(ii, c["cand_ii"]) and np.array_equal(jj, c["cand_jj"])
        ei, ej = oracle.proximity_edges(c["d"], ii, jj, cc, c["ex"][:, 0], c["ex"][:, 1], c["t0"], c["t1"], c["t"],
                                        c["rad"], c["nms"], c["thresh"], c["stereo"], c["max_factors"])
        assert np.array_equal(ei, c["es_ii"]) and np.array_equal(ej, c["es_jj"])
        fi, fj = oracle.filter_repeated_edges(ei, ej, c["act"][:, 0], c["act"][:, 1])
        assert np.array_equal(fi, c["f_ii"]) and np.array_equal(fj, c["f_jj"])
        n += 1
    assert n >= 20


def _run_device(graph, dev, c, d, cand_ii, cand_jj, cc):
    T = lambda a, dt=torch.int64: torch.as_tensor(np.ascontiguousarray(a), dtype=dt, device=dev)
    es_ii, es_jj, n_es = graph.select_edges(T(d, torch.float32), None, T(cand_ii), T(cand_jj), cc, T(c["ex"][:, 0]),
                                            T(c["ex"][:, 1]), c["t0"], c["t1"], c["t"], c["rad"], c["nms"], c["thresh"],
                                            c["stereo"], c["max_factors"])
    f_ii, f_jj, n_f = graph.filter_repeated_edges(es_ii, es_jj, n_es, T(c["act"][:, 0]), T(c["act"][:, 1]))
    n, m = int(n_es.item()), int(n_f.item())
    return es_ii[:n].cpu().numpy(), es_jj[:n].cpu().numpy(), f_ii[:m].cpu().numpy(), f_jj[:m].cpu().numpy()


@pytest.mark.gpu
def test_device_selection_matches_reference_golden(dev, db):
    from dbaf_b200 import graph
    for c in _cases():
        ci, cj, cc = graph.proximity_candidates(c["t0"], c["t1"], c["t"], c["skip"], c["window"], dev)
        assert np.array_equal(ci.cpu().numpy(), c["cand_ii"]) and np.array_equal(cj.cpu().numpy(), c["cand_jj"])
        ei, ej, fi, fj = _run_device(graph, dev, c, c["d"], c["cand_ii"], c["cand_jj"], cc)
        assert np.array_equal(ei, c["es_ii"]) and np.array_equal(ej, c["es_jj"])
        assert np.array_equal(fi, c["f_ii"]) and np.array_equal(fj, c["f_jj"])


@pytest.mark.gpu
def test_device_selection_matches_oracle_on_random_windows(dev, db, oracle):
    """40 more windows, including NaN / zero / tied distances, empty graphs and a larger initialisation grid."""
    from dbaf_b200 import graph
    rng = np.random.default_rng(77)
    for trial in range(40):
        big = trial % 8 == 7
        t = int(rng.integers(30, 70)) if big else int(rng.integers(3, 30))
        window = int(rng.choice([5, 8, 12, 25]))
        t0, t1 = (0, 0) if big or trial % 5 == 0 else (max(t - 5, 0), max(t - window, 0))
        n_ex = int(rng.integers(0, 60))
        c = dict(t=t, t0=t0, t1=t1, rad=int(rng.integers(0, 4)), nms=int(rng.integers(0, 4)),
                 thresh=float(rng.choice([8.0, 16.0, 30.0])), stereo=bool(rng.random() < 0.3),
                 max_factors=int(rng.choice([0, 10, 48, 96, 1000])), window=window,
                 skip=[[], [-4, -5, -6], [-1]][int(rng.integers(0, 3))],
                 ex=rng.integers(0, t, (n_ex, 2)), act=rng.integers(0, t, (int(rng.integers(0, 40)), 2)))
        ii, jj, cc = oracle.proximity_candidates(t0, t1, t, c["skip"], window)
        d = rng.gamma(2.0, 6.0, len(ii)).astype(np.float32)
        d[rng.random(len(ii)) < 0.05] = 500.0
        if trial % 3 == 0:
            d[rng.random(len(ii)) < 0.05] = np.nan
            d = np.round(d)                                 # many exact ties: the lower index must win
        want = oracle.proximity_edges(d, ii, jj, cc, c["ex"][:, 0], c["ex"][:, 1], t0, t1, t, c["rad"], c["nms"],
                                      c["thresh"], c["stereo"], c["max_factors"])
        wf = oracle.filter_repeated_edges(want[0], want[1], c["act"][:, 0], c["act"][:, 1])
        ei, ej, fi, fj = _run_device(graph, dev, c, d, ii, jj, cc)
        assert np.array_equal(ei, want[0]) and np.array_equal(ej, want[1]), trial
        assert np.array_equal(fi, wf[0]) and np.array_equal(fj, wf[1]), trial


@pytest.mark.gpu
def test_proximity_factors_end_to_end(dev, db, oracle, synth):
    """From the DepthVideo buffers to the edge list in one call: frame distances by k_frame_distance (both
    directions), selection and duplicate filter on the device, against the restatement fed with the same distances."""
    from dbaf_b200 import graph
    w = synth.make_window("c2", seed=3)
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    poses, disps, intr = T(w["poses"]), T(w["disps"]), T(w["intrinsics"])
    t = w["poses"].shape[0]
    t0, t1 = t - 5, max(t - 12, 0)
    act = (T(w["ii"][:20]), T(w["jj"][:20]))
    bad = (T(w["ii"][20:24]), T(w["jj"][20:24]))
    inac = (T(w["ii"][24:30]), T(w["jj"][24:30]))
    kw = dict(rad=2, nms=1, beta=0.3, thresh=60.0, stereo=False, max_factors=48, skip_edge=[-4, -5, -6], frontend_window=5)
    gi, gj = graph.proximity_factors(poses, disps, intr, t, t0, t1, act, bad, inac, **kw)
    ci, cj, cc = oracle.proximity_candidates(t0, t1, t, kw["skip_edge"], kw["frontend_window"])
    d1 = db.frame_distance(poses, disps, intr, T(ci), T(cj), 0.3)
    d2 = db.frame_distance(poses, disps, intr, T(cj), T(ci), 0.3)
    d = (0.5 * (d1 + d2)).cpu().numpy()
    ex_i = np.concatenate([w["ii"][:30]])
    ex_j = np.concatenate([w["jj"][:30]])
    ei, ej = oracle.proximity_edges(d, ci, cj, cc, ex_i, ex_j, t0, t1, t, 2, 1, 60.0, False, 48)
    keep_i = np.concatenate([w["ii"][:20], w["ii"][24:30]])
    keep_j = np.concatenate([w["jj"][:20], w["jj"][24:30]])
    fi, fj = oracle.filter_repeated_edges(ei, ej, keep_i, keep_j)
    assert len(fi) > 0 and np.isfinite(d).any()
    assert np.array_equal(gi.cpu().numpy(), fi) and np.array_equal(gj.cpu().numpy(), fj)
