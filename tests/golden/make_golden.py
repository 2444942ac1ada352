"""Generates tests/golden/ref_<config>.npz by running the UNMODIFIED reference CUDA extension
(oracle/_ref/droid_backends_ref.so, built by oracle/build_ref.py) plus the exact torch calls of
dbaf/modules/corr.py on a GPU.  Run on the B200 box:

    python tests/golden/make_golden.py gpurun_out/golden

and copy the .npz files into tests/golden/.  Inputs are not stored: they are regenerated from
the seeds below by dbaf_b200.synth (pure numpy, deterministic); a checksum of the inputs is stored
so that a drifting generator is detected.
"""
import hashlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "dba-fusion_b200"), os.path.join(ROOT, "oracle")]

from dbaf_b200 import synth  # noqa: E402

GOLDEN_CASES = {
    "tiny": dict(config="tiny", seed=0, sensor_frac=0.3, near_frac=0.02, stereo_edges=1),
    "c1": dict(config="c1", seed=0, sensor_frac=0.0, near_frac=0.0, stereo_edges=0),
}
N_CORR_EDGES = 3


def checksum(win):
    h = hashlib.sha256()
    for k in ("poses", "disps", "disps_sens", "intrinsics", "targets", "weights", "eta", "ii", "jj"):
        h.update(np.ascontiguousarray(win[k]).tobytes())
    return h.hexdigest()


def corr_inputs(case, win):
    feats = synth.make_features(case["config"], seed=case["seed"], n_frames=win["poses"].shape[0])
    coords = synth.lookup_coords(win, oob_frac=0.05, seed=case["seed"])[:N_CORR_EDGES]
    ii, jj = win["ii"][:N_CORR_EDGES], win["jj"][:N_CORR_EDGES]
    return feats[ii], feats[jj], coords


def main(outdir):
    import build_ref
    ref = build_ref.load()
    assert ref is not None, "build oracle/_ref first"
    dev = torch.device("cuda:0")
    os.makedirs(outdir, exist_ok=True)
    for name, case in GOLDEN_CASES.items():
        win = synth.make_window(**case)
        T = {k: (torch.from_numpy(v).to(dev) if isinstance(v, np.ndarray) else v) for k, v in win.items()}
        out = {"checksum": np.frombuffer(bytes.fromhex(checksum(win)), np.uint8)}
        t0, t1 = win["t0"], win["t1"]

        def fresh():
            return T["poses"].clone(), T["disps"].clone()

        common = (T["intrinsics"], T["disps_sens"], T["targets"], T["weights"], T["eta"], T["ii"], T["jj"])
        # ba, 2 full iterations
        p, d = fresh()
        dx, dz = ref.ba(p, d, *common, t0, t1, 2, 1e-4, 0.1, False)
        out.update(ba_poses=p.cpu().numpy(), ba_disps=d.cpu().numpy(), ba_dx=dx.cpu().numpy(), ba_dz=dz.cpu().numpy())
        # ba, 1 iteration (single-step comparison is better conditioned)
        p, d = fresh()
        dx, dz = ref.ba(p, d, *common, t0, t1, 1, 1e-4, 0.1, False)
        out.update(ba1_poses=p.cpu().numpy(), ba1_disps=d.cpu().numpy(), ba1_dx=dx.cpu().numpy(), ba1_dz=dz.cpu().numpy())
        # motion only
        p, d = fresh()
        dx, _ = ref.ba(p, d, *common, t0, t1, 2, 1e-4, 0.1, True)
        out.update(mo_poses=p.cpu().numpy(), mo_dx=dx.cpu().numpy())
        # BACore: hessian export + retract with a fixed dx
        p, d = fresh()
        core = ref.BACore()
        core.init(p, d, *common, t0, t1, 2, 1e-4, 0.1, False)
        n = 6 * (t1 - t0)
        H = torch.zeros(n, n, dtype=torch.float64)
        v = torch.zeros(n, dtype=torch.float64)
        core.hessian(H, v)
        rng = np.random.default_rng(7)
        dxe = torch.from_numpy(rng.normal(0, 1e-3, n))
        dxr, dzr = core.retract(dxe)
        out.update(core_H=H.numpy(), core_v=v.numpy(), core_dx_in=dxe.numpy(), core_dx=dxr.cpu().numpy(),
                   core_dz=dzr.cpu().numpy(), core_poses=p.cpu().numpy(), core_disps=d.cpu().numpy())
        del core
        # geometry
        out["frame_distance"] = ref.frame_distance(T["poses"], T["disps"], T["intrinsics"], T["ii"], T["jj"], 0.3).cpu().numpy()
        c, va = ref.projmap(T["poses"], T["disps"], T["intrinsics"], T["ii"][:4].contiguous(), T["jj"][:4].contiguous())
        out.update(projmap_coords=c.cpu().numpy(), projmap_valid=va.cpu().numpy())
        out["iproj"] = ref.iproj(T["poses"], T["disps"], T["intrinsics"]).cpu().numpy()[:3]
        ix = torch.arange(min(4, win["poses"].shape[0]), device=dev)
        th = torch.full((len(ix),), 0.05, device=dev)
        out["depth_filter"] = ref.depth_filter(T["poses"], T["disps"], T["intrinsics"], ix, th).cpu().numpy()
        if name != "tiny":
            np.savez_compressed(os.path.join(outdir, f"ref_{name}.npz"), **out)
            continue
        # correlation: exactly CorrBlock (dbaf/modules/corr.py:24-50)
        f1, f2, coords = corr_inputs(case, win)
        f1t, f2t = torch.from_numpy(f1).to(dev)[None], torch.from_numpy(f2).to(dev)[None]
        batch, num, dim, ht, wd = f1t.shape
        a = f1t.reshape(batch * num, dim, ht * wd) / 4.0
        b = f2t.reshape(batch * num, dim, ht * wd) / 4.0
        corr = torch.matmul(a.transpose(1, 2), b).view(batch, num, ht, wd, ht, wd)
        corr = corr.reshape(batch * num * ht * wd, 1, ht, wd)
        ct = torch.from_numpy(coords).to(dev)
        for l in range(4):
            vol = corr.view(batch * num, ht, wd, ht // 2 ** l, wd // 2 ** l)
            if l == 0:
                out["vol0_sample"] = vol[:, ::3, ::3].cpu().numpy()
            else:
                out[f"vol{l}"] = vol.cpu().numpy()
            look, = ref.corr_index_forward(vol.contiguous(), (ct / 2 ** l).contiguous(), 3)
            out[f"lookup{l}"] = look.cpu().numpy()
            corr = torch.nn.functional.avg_pool2d(corr, 2, stride=2)
        # fp32 lookup + backward on level 2
        vol2 = torch.from_numpy(out["vol2"]).to(dev).float().contiguous()
        look32, = ref.corr_index_forward(vol2, (ct / 4).contiguous(), 3)
        out["lookup2_f32"] = look32.cpu().numpy()
        g = torch.from_numpy(np.random.default_rng(3).normal(0, 1, look32.shape).astype(np.float32)).to(dev)
        out["lookup2_bwd"] = ref.corr_index_backward(vol2, (ct / 4).contiguous(), g, 3)[0].cpu().numpy()
        # altcorr (fp32, pixel-major features, as AltCorrBlock prepares them)
        af1 = (torch.from_numpy(f1).to(dev).float() / 4.0).permute(0, 2, 3, 1).contiguous()
        af2 = (torch.from_numpy(f2).to(dev).float() / 4.0).permute(0, 2, 3, 1).contiguous()
        ac = ct.permute(0, 2, 3, 1)[:, None].contiguous()
        alt, = ref.altcorr_forward(af1, af2, ac, 3)
        out["altcorr"] = alt.cpu().numpy()
        ga = torch.from_numpy(np.random.default_rng(4).normal(0, 1, alt.shape).astype(np.float32)).to(dev)
        g1, g2, gc = ref.altcorr_backward(af1, af2, ac, ga, 3)
        out.update(altcorr_g1=g1[:1].cpu().numpy(), altcorr_g2=g2[:1].cpu().numpy())
        torch.cuda.synchronize()
        path = os.path.join(outdir, f"ref_{name}.npz")
        np.savez_compressed(path, **out)
        print(name, "->", path, f"{os.path.getsize(path) / 1e6:.2f} MB")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
