"""Run the UNMODIFIED reference optimiser (`reconstruct/optimizer.py:90-205`, imported through ref_harness) on
synthetic detections and time it.  TEST INFRASTRUCTURE: the reference arm of `bench.py` (`--impl reference`, the
`cpu_baseline` and `reference_gpu` legs) runs this file in a child process; nothing in the product imports it.

    python oracle/ref_runner.py --device cpu|cuda --n N --M 2048 --n-bg 200 [--iters 10] [--threads T] [--warm 1]

prints one JSON object: seconds per detection, objects/s, threads, torch version, where the reference came from."""
import argparse
import contextlib
import io
import json
import os
import sys
import tempfile
import time
import warnings

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


class Reference:
    def __init__(self, device="cpu", threads=None):
        sys.path.insert(0, HERE)
        import ref_harness
        self.h = ref_harness
        self.ref = ref_harness.import_reference(force_cpu=(device == "cpu"))
        sys.path.append(os.path.join(ROOT, "dsp-slam-rgbd_b200"))
        import torch
        from dspslam import synth           # numpy generators; `reconstruct` / `deep_sdf` resolve to the reference
        self.torch, self.synth = torch, synth
        warnings.simplefilter("ignore")
        if threads:
            torch.set_num_threads(int(threads))
        self.threads = torch.get_num_threads()
        self.device = device

    def decoder(self):
        """The W-fit synthetic decoder, loaded through the reference's own get_decoder (utils.py / workspace.py:202-223)."""
        import numpy as np
        torch, synth, ref = self.torch, self.synth, self.ref
        with np.load(synth.WFIT_PATH) as z:
            sd = {k: torch.from_numpy(z[k].astype(np.float32)) for k in z.files}
        tmp = tempfile.mkdtemp()
        dec = ref.decoder.Decoder(synth.CODE_LEN, **synth.NETWORK_SPECS)
        dec.load_state_dict(sd)
        synth.write_experiment_dir(dec, tmp)
        d = ref.utils.get_decoder(ref.utils.ForceKeyErrorDict(DeepSDF_DIR=tmp))
        for p in d.parameters():
            p.requires_grad_(False)
        return d

    def optimizer(self, iters=None):
        cfg = self.ref.utils.get_configs(os.path.join(self.h.REF_ROOT, "configs", "config_kitti.json"))
        if iters is not None:
            cfg.optimizer.joint_optim.num_iterations = int(iters)
        return self.ref.optimizer.Optimizer(self.decoder(), cfg)

    def reconstruct(self, opt, det):
        with contextlib.redirect_stdout(io.StringIO()):
            out = opt.reconstruct_object(det["t_cam_obj"].copy(), det["pts"], det["rays"], det["depth"])
        if self.device == "cuda":
            self.torch.cuda.synchronize()
        return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--device", default="cpu", choices=["cpu", "cuda"])
    ap.add_argument("--n", type=int, default=1)
    ap.add_argument("--first", type=int, default=0)
    ap.add_argument("--M", type=int, default=2048)
    ap.add_argument("--n-bg", type=int, default=200)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--warm", type=int, default=1)
    a = ap.parse_args()
    R = Reference(a.device, a.threads or (os.cpu_count() if a.device == "cpu" else None))
    opt = R.optimizer(a.iters)
    for k in range(a.warm):                                  # page-in, thread pools, cuDNN/cuBLAS handles
        R.reconstruct(opt, R.synth.make_kitti_detection(10_000 + k, M=250, n_bg=200))
    dets = [R.synth.make_kitti_detection(a.first + k, M=a.M, n_bg=a.n_bg) for k in range(a.n)]
    secs, good = [], 0
    for d in dets:
        t0 = time.perf_counter()
        out = R.reconstruct(opt, d)
        secs.append(time.perf_counter() - t0)
        good += bool(out.is_good)
    tot = sum(secs)
    print(json.dumps({"device": a.device, "n": a.n, "good": good, "s_per_object": secs, "objects_per_s": a.n / tot,
                      "threads": R.threads, "torch": R.torch.__version__, "ref_root": R.h.REF_ROOT,
                      "M": a.M, "N": a.M + a.n_bg, "iters": a.iters}), flush=True)


if __name__ == "__main__":
    main()
