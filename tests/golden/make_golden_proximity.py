"""Golden vectors for the frame-graph edge selection from the reference's own Python
(dbaf/covisible_graph.py:61-72 `__filter_repeated_edges`, :357-441 `add_proximity_factors`).

The module cannot be imported here (lietorch, cv2, matplotlib), so the two methods are cut out of the file with
`ast` at generation time and executed unmodified against a stand-in `self` whose `video.distance` returns a
prepared float32 vector.  Nothing of the reference is stored: only the inputs and the edge lists it produced.

usage: python tests/golden/make_golden_proximity.py   (needs /root/reference; writes tests/golden/ref_proximity.npz)
"""
import ast
import os
import textwrap
import types

import numpy as np
import torch

REF = os.environ.get("DBAF_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_proximity.npz")


def reference_methods():
    src = open(os.path.join(REF, "dbaf", "covisible_graph.py")).read()
    tree = ast.parse(src)
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "CovisibleGraph")
    ns = dict(torch=torch, np=np)
    for fn in cls.body:
        if isinstance(fn, ast.FunctionDef) and fn.name in ("add_proximity_factors", "__filter_repeated_edges"):
            code = textwrap.dedent(ast.get_source_segment(src, fn))
            code = code.replace("def __filter_repeated_edges", "def filter_repeated_edges")
            exec(compile(code, fn.name, "exec"), ns)
    return ns["add_proximity_factors"], ns["filter_repeated_edges"]


def make_case(rng, kind):
    if kind == "frontend":                      # dbaf_frontend.py:241: t0 = t1'-5, t1 = max(t1'-window, 0)
        window = int(rng.choice([5, 8, 12, 25]))
        t = int(rng.integers(6, 40))
        t0, t1 = max(t - 5, 0), max(t - window, 0)
    else:                                       # dbaf_frontend.py:830: initialisation, whole buffer
        window = int(rng.choice([5, 8]))
        t = int(rng.integers(4, 14))
        t0 = t1 = 0
    rad = int(rng.choice([1, 2, 3]))
    nms = int(rng.choice([0, 1, 2]))
    thresh = float(rng.choice([12.0, 16.0, 24.0]))
    stereo = bool(rng.random() < 0.3)
    max_factors = int(rng.choice([16, 24, 48, 200]))
    skip = [[], [-4, -5, -6], [-2, -3]][int(rng.integers(0, 3))]
    n_ex = int(rng.integers(0, 40))
    ex = rng.integers(max(t1 - 2, 0), t, (n_ex, 2))
    return dict(t=t, t0=t0, t1=t1, rad=rad, nms=nms, thresh=thresh, stereo=stereo, max_factors=max_factors,
                skip=skip, window=window, ex=ex, seed=int(rng.integers(0, 1 << 30)))


def run_reference(add_prox, filt, c):
    t, t0, t1 = c["t"], c["t0"], c["t1"]
    drng = np.random.default_rng(c["seed"])
    dist = {}

    def distance(ii, jj, beta=0.3):
        d = drng.gamma(2.0, 6.0, len(ii)).astype(np.float32)       # mostly below the thresholds, some above 100
        d[drng.random(len(ii)) < 0.05] = 1000.0
        # exact zeros only among the skip entries (they exercise the "0 < d" test of the tail rule): equal distances
        # inside the grid would leave the order to torch.argsort, which does not specify it for ties
        cc = (t - t0) * (t - t1)
        z = drng.random(len(ii)) < 0.3
        z[:cc] = False
        d[z] = 0.0
        dist["d"], dist["ii"], dist["jj"] = d.copy(), ii.numpy().copy(), jj.numpy().copy()
        return torch.from_numpy(d)

    got = {}
    ex = c["ex"]
    n1, n2 = len(ex) // 3, 2 * len(ex) // 3
    as_t = lambda a: torch.as_tensor(a, dtype=torch.long)
    self = types.SimpleNamespace(
        video=types.SimpleNamespace(counter=types.SimpleNamespace(value=t), distance=distance, stereo=c["stereo"]),
        skip_edge=c["skip"], frontend_window=c["window"], max_factors=c["max_factors"], device="cpu",
        ii=as_t(ex[:n1, 0]), jj=as_t(ex[:n1, 1]), ii_bad=as_t(ex[n1:n2, 0]), jj_bad=as_t(ex[n1:n2, 1]),
        ii_inac=as_t(ex[n2:, 0]), jj_inac=as_t(ex[n2:, 1]),
        add_factors=lambda ii, jj, remove=False: got.update(ii=ii.numpy().copy(), jj=jj.numpy().copy()))
    add_prox(self, t0=t0, t1=t1, rad=c["rad"], nms=c["nms"], beta=0.25, thresh=c["thresh"], remove=False)
    fi, fj = filt(self, torch.from_numpy(got["ii"]), torch.from_numpy(got["jj"]))
    act = np.concatenate([ex[:n1], ex[n2:]])                       # active + inactive: what the filter looks at
    return dist, got, (fi.numpy(), fj.numpy()), act


def main():
    add_prox, filt = reference_methods()
    rng = np.random.default_rng(2024)
    out = {}
    n = 0
    for kind in ["frontend"] * 24 + ["init"] * 8:
        c = make_case(rng, kind)
        try:
            dist, got, filt_out, act = run_reference(add_prox, filt, c)
        except Exception as ex:                                     # e.g. an empty grid: torch.max of nothing
            print("skipped a case:", type(ex).__name__, ex)
            continue
        p = f"case{n}_"
        out[p + "params"] = np.array([c["t"], c["t0"], c["t1"], c["rad"], c["nms"], int(c["stereo"]), c["max_factors"],
                                      c["window"]], np.int64)
        out[p + "thresh"] = np.float32(c["thresh"])
        out[p + "skip"] = np.asarray(c["skip"], np.int64)
        out[p + "ex"] = c["ex"].astype(np.int64)
        out[p + "act"] = act.astype(np.int64)
        out[p + "d"], out[p + "cand_ii"], out[p + "cand_jj"] = dist["d"], dist["ii"], dist["jj"]
        out[p + "es_ii"], out[p + "es_jj"] = got["ii"].astype(np.int64), got["jj"].astype(np.int64)
        out[p + "f_ii"], out[p + "f_jj"] = filt_out[0].astype(np.int64), filt_out[1].astype(np.int64)
        n += 1
    out["n_cases"] = np.int64(n)
    np.savez_compressed(OUT, **out)
    print(f"wrote {OUT}: {n} cases")


if __name__ == "__main__":
    main()
