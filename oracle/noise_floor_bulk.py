"""Noise floor of the reference's free-running END state over the whole C1 population.  TEST INFRASTRUCTURE
(build container only: imports /root/reference).      Usage: python oracle/noise_floor_bulk.py [n] [which]

oracle/noise_floor.py showed on 4 objects that the reference does not reproduce its own end state when its FP32
rounding changes.  This script puts a population number on it, in north_star's own units: the UNMODIFIED reference
`reconstruct_object` on the objects of tests/golden/bulk_c1.npz (or bulk_c1_heldout.npz), inputs identical except that
the surface points are moved by ONE FP32 ulp, end state against the end state stored in the golden file (the same
code, same thread count, unperturbed).  "within" = 1 cm, 0.1 deg, |dcode| / max(|code|, 0.1) < 1e-2 - exactly what
tools/parity_stats.py: free_run_stats applies to the library.  The fraction is the ceiling any other FP32-grade
implementation can be held to for a free run; tests/test_gpu_parity_bulk.py compares the library's modes with it.
Writes tests/golden/noise_floor_bulk[_heldout].json.
"""
import contextlib
import io
import json
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
import make_goldens as mg  # noqa: E402  (imports the reference through ref_harness)

warnings.simplefilter("ignore")
synth = mg.synth


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
    which = sys.argv[2] if len(sys.argv) > 2 else "bulk_c1.npz"
    with np.load(os.path.join(ROOT, "tests", "golden", which)) as z:
        b = {k: z[k] for k in z.files}
    k0 = int(b["k0"]) if "k0" in b else 0
    with np.load(synth.WFIT_PATH) as z:
        wfit = {k: mg.torch.from_numpy(z[k].astype(np.float32)) for k in z.files}
    opt = mg.ref.optimizer.Optimizer(mg.ref_decoder(wfit), mg.ref_config("kitti"))
    rows = []
    for k in range(min(n, b["T_oc"].shape[0])):
        if not b["is_good"][k]:
            continue
        d = synth.make_kitti_detection(k0 + k)
        pts = np.nextafter(d["pts"].astype(np.float32), np.float32(np.inf)).astype(np.float32)
        with contextlib.redirect_stdout(io.StringIO()):
            out = opt.reconstruct_object(d["t_cam_obj"].copy(), pts, d["rays"].copy(), d["depth"].copy())
        if not out.is_good:
            rows.append(dict(k=k, dt_m=1.0, rot_deg=90.0, code_rel=1.0))
            continue
        dt, ang, _ = synth.pose_errors(np.asarray(out.t_cam_obj, np.float32), b["t_cam_obj"][k])
        cn = b["code"][k]
        rows.append(dict(k=k, dt_m=float(dt), rot_deg=float(ang),
                         code_rel=float(np.linalg.norm(np.asarray(out.code, np.float32) - cn) / max(np.linalg.norm(cn), 0.1))))
        if k % 8 == 0:
            print(rows[-1], flush=True)
    E = np.array([[r["dt_m"], r["rot_deg"], r["code_rel"]] for r in rows])
    ok = (E[:, 0] < 1e-2) & (E[:, 1] < 0.1) & (E[:, 2] < 1e-2)
    pose_ok = (E[:, 0] < 1e-2) & (E[:, 1] < 0.1)
    q = lambda a, p: float(np.quantile(a, p))
    summary = {"golden": which, "variant": "surface points + 1 FP32 ulp, otherwise the run that made the golden file",
               "objects": len(rows), "within": float(ok.mean()), "within_pose_only": float(pose_ok.mean()),
               "dt_m": {"median": q(E[:, 0], .5), "q95": q(E[:, 0], .95)},
               "rot_deg": {"median": q(E[:, 1], .5), "q95": q(E[:, 1], .95)},
               "code_rel": {"median": q(E[:, 2], .5), "q95": q(E[:, 2], .95)}}
    name = "noise_floor_bulk.json" if "heldout" not in which else "noise_floor_bulk_heldout.json"
    with open(os.path.join(ROOT, "tests", "golden", name), "w") as f:
        json.dump({"summary": summary, "objects": rows}, f, indent=1)
    print(json.dumps(summary))


if __name__ == "__main__":
    main()
