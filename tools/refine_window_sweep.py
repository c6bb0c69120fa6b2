"""Does the FP16 rounding of the ray-sample SDF VALUES (not only the decisions at |sdf| = cut_off) move the free-running
end state?  Sweeps joint_params.refine_window: a sample whose first-pass |sdf| lies within window * cut_off of the
threshold is evaluated a second time at FP32 grade; window 1.0 covers the whole occupancy band (0 < |sdf| < 2 cut_off),
i.e. every sample whose value enters a residual.  Prints, per (mode, window): teacher-forced and free-running
fractions on bulk_c1 / held-out (tools/parity_stats.py) and the C2 throughput (resident batch, 5 steps).
Usage: python tools/refine_window_sweep.py [mode ...]"""
import json
import os
import sys

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path[:0] = [os.path.join(ROOT, "dsp-slam-rgbd_b200"), ROOT, os.path.join(ROOT, "tools")]
import parity_stats  # noqa: E402
import bench  # noqa: E402


def main():
    import torch
    from dspslam import synth
    from dspslam.engine import Engine
    eng = Engine(synth.load_wfit())
    batch = eng.upload_joint(eng.pack_joint(bench.make_dets(0)))
    for mode in sys.argv[1:] or ["tc", "tc_exact"]:
        for w in (0.05, 0.25, 0.5, 1.0):
            tag = "%s:w%g" % (mode, w)
            row = {"mode": mode, "refine_window": w}
            for name, which in (("c1", "bulk_c1.npz"), ("heldout", "bulk_c1_heldout.npz")):
                st = parity_stats.step_stats(eng, tag, 128, which)
                fr = parity_stats.free_run_stats(eng, tag, 128, which)
                row[name] = {"steps": round(st["steps_within"], 4), "last": round(st["objects_last_step_within"], 4),
                             "all_steps": round(st["objects_all_steps_within"], 4), "free_run": round(fr["within"], 4),
                             "free_run_loose": round(fr["within_5cm_1deg_1e-1"], 4),
                             "free_code_rel_median": round(fr["code_rel"]["median"], 5)}
            prm = parity_stats.joint_params(synth.kitti_config(), 10)
            prm["refine_window"] = w
            for _ in range(2):
                eng.run_joint(batch, prm, mode=mode)
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
            torch.cuda.synchronize()
            ev[0].record()
            for _ in range(5):
                eng.run_joint(batch, prm, mode=mode)
            ev[1].record()
            torch.cuda.synchronize()
            row["c2_objects_per_s"] = round(16 * 5 / (ev[0].elapsed_time(ev[1]) * 1e-3), 1)
            print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
