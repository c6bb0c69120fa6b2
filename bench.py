#!/usr/bin/env python
"""Benchmark of the B200 DSP-SLAM object-reconstruction hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--mode tc|fp32]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Workload (`config.workload`): BASELINE.json configs[1] = "C2", a KITTI stereo+LiDAR frame batch of 16 car
detections x 2048 surface points + 2248 rays x 50 depth samples, 10 Gauss-Newton iterations
(SURVEY.md 8(d)), synthetic detections and the W-fit synthetic decoder (assets/).  One STEP = one pass of
the hot path (`dspslam_gn_joint`: all 10 iterations) over one batch of 16 detections per GPU.  With N GPUs
every rank optimises its own 16 detections (independent objects, no collective: weak scaling).

  value     objects/s with the batch already resident in HBM (device-timed with CUDA events, max over ranks)
  e2e       objects/s through the reference-shaped API `Optimizer.reconstruct_objects` with HOST numpy
            arrays: packing, pinned H2D, the GN loop and the D2H of poses/codes inside the timed region
  roofline  the dominant kernel (forward-only decoder over the in-ball ray samples of the render term):
            algorithmic FLOP (3 671 040 per query) / its CUDA-event time measured live inside the timed
            region via dspslam_profile_*; peak = MEASURED_PEAKS.json bf16 dense (sustained figure: the
            kernel is timed inside a long step).
  cpu_baseline  the UNMODIFIED reference optimiser (oracle/_ref: verbatim copy of the reference's hot-path
            files made by oracle/fetch_ref.py; git-ignored, ships to the box like the .so) timed on this box's
            host cores on a bounded sample of the same workload (rank 0, N = 1 only); `reference_gpu` is the
            same unmodified code with its tensors on the B200 (its own .cuda() path, stock PyTorch kernels).

`--impl reference` times that CPU implementation alone, one C2 detection per step, and prints the same JSON
line (falls back to the numpy oracle port, kind "port", only if oracle/_ref is absent).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "dsp-slam-rgbd_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

FLOP_FWD = 3_671_040          # forward-only SDF query  
FLOP_JAC = 7_342_080          # SDF + Jacobian query
OBJ_PER_RANK = 16
M_PTS, N_BG, D_SAMPLES, ITERS = 2048, 200, 50, 10
METRIC = "objects_reconstructed_per_s"


def workload_desc():
    return ("C2: KITTI stereo+LiDAR frame batch, %d car detections/GPU x %d surface pts + %d rays x %d depth "
            "samples, %d GN iters" % (OBJ_PER_RANK, M_PTS, M_PTS + N_BG, D_SAMPLES, ITERS))


def make_dets(rank, n=OBJ_PER_RANK):
    from dspslam import synth
    return [synth.make_kitti_detection(rank * OBJ_PER_RANK + k, M=M_PTS, n_bg=N_BG) for k in range(n)]


KERNEL_SOURCES = ("kernels_tc.cu", "kernels_tc.h", "tc_ptx.cuh", "common.cuh", "gn_device.cuh")


def kernel_src_sha16():
    """sha256 (first 16 hex) over the sources of the tcgen05 kernels: ties a committed ncu capture to a build."""
    import hashlib
    h = hashlib.sha256()
    for n in KERNEL_SOURCES:
        with open(os.path.join(ROOT, "dsp-slam-rgbd_b200", "csrc", n), "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def traffic_of(kernel, with_source=False):
    """dram bytes per launch of `kernel` from the newest committed ncu capture (profiles/r*_traffic.json,
    tools/ncu_traffic.py) that was taken on THESE kernel sources and THIS workload; None when the only captures are of
    another build (a stale number next to live timings is worse than null)."""
    import glob
    res = (None, None)
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_traffic.json")), reverse=True):
        try:
            with open(path) as f:
                j = json.load(f)
            if j.get("kernel_src_sha16") != kernel_src_sha16() or j.get("workload") != workload_desc():
                continue
            for name, v in j["kernels"].items():
                if kernel in name:
                    res = (int(v["traffic_bytes"]), "%s (dram__bytes_read.sum + dram__bytes_write.sum, %s)"
                           % (os.path.relpath(path, ROOT), j.get("source")))
                    break
        except (OSError, KeyError, ValueError):
            continue
        if res[0] is not None:
            break
    return res if with_source else res[0]


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        with open(path) as f:
            p = json.load(f)
        return dict(tflops=float(p.get("bf16_tflops_sustained", p.get("bf16_tflops", 1590.0))),
                    hbm=float(p.get("hbm_gbs", 6650.0)), src="measured (MEASURED_PEAKS.json, bf16 sustained)")
    return dict(tflops=1400.0, hbm=6650.0, src="fallback (B200_PROFILING.md)")


# ------------------------------------------------------------------------------------------------
# clocks sampling (nvidia-smi during the timed region)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        threading.Thread(target=self._read, daemon=True).start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1 + 0.2 and len(r) >= 7] or [r for _, r in self.rows if len(r) >= 7]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(r[0]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        pw = [float(r[2]) for r in rows]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][1]), "reasons": reasons,
                "samples": len(rows), "power_w_max": max(pw), "power_w_mean": float(np.mean(pw))}


# ------------------------------------------------------------------------------------------------
# reference arm: the UNMODIFIED reference optimiser (oracle/_ref, a verbatim copy of the reference's own
# reconstruct/*.py + deep_sdf/*.py made by oracle/fetch_ref.py) on the host cores, in a child process
# (oracle/ref_runner.py); the numpy oracle port only where that copy is absent
# ------------------------------------------------------------------------------------------------
_CAPS = ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "OPENBLAS_NUM_THREADS", "RANK", "WORLD_SIZE", "LOCAL_RANK",
         "LOCAL_WORLD_SIZE", "GROUP_RANK", "ROLE_RANK", "MASTER_ADDR", "MASTER_PORT", "TORCHELASTIC_RUN_ID")


def reference_available():
    return (os.path.isfile(os.path.join(ROOT, "oracle", "_ref", "reconstruct", "optimizer.py"))
            or os.path.isfile("/root/reference/reconstruct/optimizer.py"))


def run_reference_child(device, n, first=0, warm=1, M=M_PTS, n_bg=N_BG, iters=ITERS, timeout=1500):
    """Runs oracle/ref_runner.py in a clean child (torchrun caps OMP_NUM_THREADS at 1 for its workers; this process
    may also hold this repo's own `reconstruct` mirror, which shadows the reference's package name)."""
    env = {k: v for k, v in os.environ.items() if k not in _CAPS}
    if device == "cpu":
        env["CUDA_VISIBLE_DEVICES"] = ""     # the CPU arm must not see the GPU (DataParallel in the reference's
                                             # config_decoder moves the module to cuda:0 whenever one is visible)
    cmd = [sys.executable, os.path.join(ROOT, "oracle", "ref_runner.py"), "--device", device, "--n", str(n), "--first",
           str(first), "--M", str(M), "--n-bg", str(n_bg), "--iters", str(iters), "--warm", str(warm)]
    out = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout)
    if out.returncode != 0:
        raise RuntimeError("reference child failed: " + out.stderr[-800:])
    return json.loads(out.stdout.strip().splitlines()[-1])


def cpu_threads():
    try:
        from threadpoolctl import threadpool_info
        n = max([i.get("num_threads", 1) for i in threadpool_info()] or [1])
        return int(n)
    except Exception:
        return os.cpu_count() or 1


def run_port_sample(n_obj, iters):
    """Fallback CPU arm (no oracle/_ref): the numpy oracle port; returns seconds per detection, threads."""
    from dspslam import synth
    from oracle import dspslam_oracle as orc
    dec = synth.load_wfit()
    cfg = synth.kitti_config()
    cfg.optimizer.joint_optim.num_iterations = iters
    opt = orc.OracleOptimizer(orc.OracleDecoder(dec.state_dict()), cfg)
    secs = []
    for d in make_dets(0, n_obj):
        t0 = time.perf_counter()
        opt.reconstruct_object(d["t_cam_obj"], d["pts"], d["rays"], d["depth"])
        secs.append(time.perf_counter() - t0)
    return secs, cpu_threads()


def cpu_sample(n_obj, warm=1):
    """(seconds per C2 detection [n_obj], threads, kind) on the host cores."""
    if reference_available():
        r = run_reference_child("cpu", n_obj, warm=warm)
        return r["s_per_object"], r["threads"], "reference"
    if warm:
        run_port_sample(1, 1)
    secs, thr = run_port_sample(n_obj, ITERS)
    return secs, thr, "port"


def shared_config(mode="tc", ray_phases=0):
    """`config` is the same object in both arms (the arms differ in `impl`, not in the workload)."""
    return {"workload": workload_desc(), "mode": mode, "objects_per_gpu": OBJ_PER_RANK, "ray_phases": ray_phases,
            "decoder": "DeepSDF 8x512, code 64, W-fit synthetic weights",
            "l2": "flushed between steps (256 MiB fill, untimed); each step times one dspslam_gn_joint"}


def reference_arm(args, rank, world):
    """One STEP of this arm = ONE detection of the C2 batch (seed = step index) through the unmodified
    `Optimizer.reconstruct_object` (all 10 GN iterations, torch CPU FP32, every host thread): a bounded sample of
    the 16-detection batch - the reference itself processes detections one at a time (src/LocalMapping_util.cc:165-206),
    so objects/s is the same rate.  `ms_per_step` is the time measured per step, `objects_per_step` = 1."""
    if rank != 0:
        return
    K = max(1, args.steps)
    secs, thr, kind = cpu_sample(K, warm=1 if args.warmup > 0 else 0)
    mean_s = float(np.mean(secs))
    v = 1.0 / mean_s
    what = ("the UNMODIFIED reference (oracle/_ref = verbatim reconstruct/optimizer.py:90-205 + loss*.py + deep_sdf, torch CPU FP32)"
            if kind == "reference" else "numpy oracle port of reconstruct/optimizer.py (oracle/_ref absent)")
    sample = ("%d steps, each ONE C2 detection (M=%d, N=%d, D=%d) x all %d GN iterations; %.2f s of CPU work per step "
              "(min %.2f, max %.2f); warm-up = %d C1-sized detection(s), untimed; decoder parameters frozen "
              "(requires_grad False, oracle/ref_runner.py - less backward work than the stock script, i.e. in the reference's favour)"
              % (K, M_PTS, M_PTS + N_BG, D_SAMPLES, ITERS, mean_s, min(secs), max(secs), 1 if args.warmup > 0 else 0))
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "objects/s", "n_gpus": args.gpus,
            "steps": K, "warmup": args.warmup, "ms_per_step": 1e3 * mean_s, "objects_per_step": 1,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": shared_config(), "reference_detail": what,
            "cpu_baseline": {"value": v, "unit": "objects/s", "cores": thr, "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": "objects/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def ours_arm(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from dspslam import synth
    from reconstruct.optimizer import Optimizer

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    dec = synth.load_wfit()
    cfg = synth.kitti_config()
    opt = Optimizer(dec, cfg, mode=args.mode, device=local_rank, ray_phases=args.ray_phases)
    eng = opt.engine
    prm = opt._params()
    dets = make_dets(rank)
    batch = eng.upload_joint(eng.pack_joint(dets))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2

    # ---- device-resident timing -------------------------------------------------------------
    for _ in range(max(3, args.warmup)):
        eng.run_joint(batch, prm)
    res = eng.download_joint(batch)
    n_good = int((res.status == 0).sum())
    eng.profile_enable(True)
    sampler = ClockSampler(local_rank)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    l0 = eng.launches
    barrier()
    sampler.start()                                      # every rank samples its own board (per_rank below)
    t_wall0 = time.time()
    for a, b in ev:
        flush.fill_(1)                                   # L2 flush between steps (untimed)
        a.record()
        eng.run_joint(batch, prm)
        b.record()
    barrier()
    t_wall1 = time.time()
    clocks = sampler.stop(t_wall0, t_wall1)
    launches = eng.launches - l0
    step_ms = [a.elapsed_time(b) for a, b in ev]
    dev_ms = max_over_ranks(float(np.sum(step_ms)))
    # per-rank step time, SM clock and board power: what the max over ranks is made of (weak-scaling attribution)
    mine = [float(np.mean(step_ms)), float((clocks or {}).get("sm_mhz") or 0.0), float((clocks or {}).get("power_w_mean") or 0.0),
            float(len((clocks or {}).get("reasons") or []))]
    if world > 1:
        tl = [torch.zeros(4, dtype=torch.float64, device=dev) for _ in range(world)]
        dist.all_gather(tl, torch.tensor(mine, dtype=torch.float64, device=dev))
        per_rank = [t.tolist() for t in tl]
    else:
        per_rank = [mine]
    prof = eng.profile_read()
    eng.profile_enable(False)

    # ---- end to end through the reference-shaped API (host arrays in, host arrays out) --------
    for _ in range(2):
        opt.reconstruct_objects(dets)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        outs = opt.reconstruct_objects(dets)
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    e2e_good = sum(1 for o in outs if o.is_good)

    total_obj = OBJ_PER_RANK * world * args.steps
    value = total_obj / (dev_ms * 1e-3)
    pk = peaks()
    fwd_flop = prof["rows_gn_fwd"] * FLOP_FWD
    jac_flop = prof["rows_gn_jac"] * FLOP_JAC
    fwd_tf = fwd_flop / (prof["ms_gn_fwd"] * 1e-3) / 1e12 if prof["ms_gn_fwd"] > 0 else 0.0
    jac_tf = jac_flop / (prof["ms_gn_jac"] * 1e-3) / 1e12 if prof["ms_gn_jac"] > 0 else 0.0
    rows_fwd_all = sum_over_ranks(prof["rows_gn_fwd"])
    rows_jac_all = sum_over_ranks(prof["rows_gn_jac"])

    extra = other_configs(args, rank, world, local_rank, dec, opt, barrier, max_over_ranks, sum_over_ranks)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    line = {
        "metric": METRIC, "value": value, "unit": "objects/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None,
        "dtype": "f16 operands, f32 accumulate (tcgen05)" if args.mode != "fp32" else "f32",
        "data": "synthetic",
        "config": shared_config(args.mode, args.ray_phases),
        "e2e": {"value": total_obj / e2e_s, "unit": "objects/s", "h2d_bytes_per_step": int(batch.h2d_bytes),
                "d2h_bytes_per_step": int(batch.d2h_bytes)},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "tensor", "kernel": "forward-only decoder over in-ball ray samples (GN render term)",
                     "achieved": fwd_tf, "peak": pk["tflops"], "unit": "TFLOP/s", "frac": fwd_tf / pk["tflops"],
                     # DRAM bytes of one forward launch (read + write) from the committed `ncu --set full` capture of this
                     # command (tools/ncu_traffic.py -> profiles/rNN_traffic.json); null unless a capture of this build exists
                     "traffic": traffic_of("k_tc<3>") if args.mode != "fp32" else None,
                     "traffic_source": (traffic_of("k_tc<3>", True)[1] or "no ncu capture of this build and workload under "
                                        "profiles/ (kernel_src_sha16 %s)" % kernel_src_sha16()),
                     "peak_source": pk["src"],
                     "avg_launch_ms": prof["ms_gn_fwd"] / max(1, prof["n_gn_fwd"]),
                     "rows_per_launch": prof["rows_gn_fwd"] / max(1, prof["n_gn_fwd"]),
                     "launch_note": "two launches per GN iteration are averaged here: the large phase-0 launch (all rays, "
                                    "~2 ms) and the small phase-1 launch (rays still alive, ~0.06 ms); achieved = all rows "
                                    "/ all launch time",
                     "share_of_step": prof["ms_gn_fwd"] / max(1e-9, float(np.sum(step_ms))),
                     "rows_evaluated": int(prof["rows_gn_fwd"]), "rows_in_ball": int(prof["rows_gn_inball"]),
                     "note": "rows_evaluated < rows_in_ball: rays stop at their first sample with sdf <= -th "
                             "(exactly zero transmittance behind it; results bit-identical to evaluating all)"},
        "roofline_jacobian": {"bound": "tensor", "kernel": "fused decoder forward + input-gradient sweep + JtJ/Jtr",
                              "achieved": jac_tf, "peak": pk["tflops"], "unit": "TFLOP/s",
                              "frac": jac_tf / pk["tflops"],
                              "avg_launch_ms": prof["ms_gn_jac"] / max(1, prof["n_gn_jac"]),
                              "rows_per_launch": prof["rows_gn_jac"] / max(1, prof["n_gn_jac"]),
                              "share_of_step": prof["ms_gn_jac"] / max(1e-9, float(np.sum(step_ms)))},
        "mquery_per_s": {"forward": rows_fwd_all / (dev_ms * 1e-3) / 1e6,
                         "sdf_jacobian": rows_jac_all / (dev_ms * 1e-3) / 1e6,
                         "forward_kernel_only": prof["rows_gn_fwd"] / max(1e-9, prof["ms_gn_fwd"] * 1e-3) / 1e6,
                         "sdf_jacobian_kernel_only": prof["rows_gn_jac"] / max(1e-9, prof["ms_gn_jac"] * 1e-3) / 1e6},
        "objects_good": {"resident": n_good, "e2e": e2e_good, "of": OBJ_PER_RANK},
        "refine": {"rows": int(prof["rows_gn_refine"]), "launches": int(prof["n_gn_refine"]), "ms": prof["ms_gn_refine"],
                   "share_of_step": prof["ms_gn_refine"] / max(1e-9, float(np.sum(step_ms))),
                   "note": "ray samples within refine_window * cut_off of a decision threshold, evaluated a second "
                           "time with hi+lo FP16 operand pairs (k_tc_jacx<2>) before the per-ray scan"},
    }
    if clocks and clocks.get("power_w_mean"):
        # energy per object on rank 0's board while the timed steps ran (nvidia-smi power.draw, 100 ms samples; the
        # region also holds the untimed L2 flushes, so this is a lower bound on objects per joule)
        line["objects_per_joule"] = (value / world) / clocks["power_w_mean"]
    line.update(extra)
    if world == 1 and not args.no_cpu:
        secs, thr, kind = cpu_sample(2, warm=1)
        line["cpu_baseline"] = {"value": 1.0 / float(np.mean(secs)), "unit": "objects/s", "cores": thr, "kind": kind,
                                "sample": "2 of the %d C2 detections x all %d GN iterations, one at a time as the reference "
                                          "runs them (%.1f s of CPU work), %s" % (
                                              OBJ_PER_RANK, ITERS, sum(secs),
                                              "unmodified reference optimiser (oracle/_ref), torch CPU FP32" if kind == "reference"
                                              else "numpy oracle port")}
        if reference_available():
            # like-for-like GPU figure: the unmodified reference on this B200 through its own .cuda() path
            try:
                r = run_reference_child("cuda", 4, warm=2, timeout=600)
                line["reference_gpu"] = {"value": r["objects_per_s"], "unit": "objects/s", "kind": "reference",
                                         "sample": "4 C2 detections, one at a time, unmodified reference with every tensor "
                                                   "on this GPU (stock PyTorch %s FP32 kernels)" % r["torch"]}
            except Exception as ex:  # noqa: BLE001
                line["reference_gpu"] = {"unavailable": str(ex)[-300:]}
    else:
        line["cpu_baseline"] = None
    # compact, numbers-only digest as the LAST key of the line: a record that keeps only the tail of the line still
    # holds every configuration (c1..c5), the operating points, the per-rank split and the parity fractions
    r3 = lambda x: None if x is None else round(float(x), 3)
    par = extra.get("parity") or {}
    frac = lambda k, f: r3((par.get(k) or {}).get(f))
    tail = {"n_gpus": world, "mode": args.mode, "objects_per_s": r3(value), "e2e_objects_per_s": r3(line["e2e"]["value"]),
            "wsplit_mask": int(eng.lib.dspslam_query(eng._h, 0)), "jac_early": int(eng.lib.dspslam_query(eng._h, 1)),
            "per_rank": {"ms_per_step": [r3(r[0]) for r in per_rank], "sm_mhz": [r3(r[1]) for r in per_rank],
                         "power_w": [r3(r[2]) for r in per_rank], "throttle_reasons": [int(r[3]) for r in per_rank]}}
    if "c1" in extra:
        tail["c1_ms"] = r3(extra["c1"]["c1_ms_per_object"])
    if "c3" in extra:
        tail["c3_objects_per_s"] = r3(extra["c3"]["objects_per_s"])
    if "c4" in extra:
        c4 = extra["c4"]
        tail["c4"] = {k: r3(c4[k]) for k in ("ms_grid", "ms_slab_compute", "ms_allgather_exposed", "ms_allgather_alone",
                                             "tflops", "hbm_gbs_algorithmic")}
        tail["c4"]["pieces"] = c4["pieces"]
    if "c5" in extra:
        tail["c5_objects_per_s"] = r3(extra["c5"]["objects_per_s"])
    if "modes_objects_per_s_per_gpu" in extra:
        tail["modes"] = {k: r3(v) for k, v in extra["modes_objects_per_s_per_gpu"].items() if k != "note"}
    if par and "error" not in par:
        tail["parity"] = {"jac_rows": frac("jacobian", "rows_within"), "c1_steps": frac("c1", "steps_within"),
                          "c1_last": frac("c1", "objects_last_step_within"), "c1_free_run": frac("c1_free_run", "within"),
                          "heldout_steps": frac("c1_heldout", "steps_within"),
                          "heldout_last": frac("c1_heldout", "objects_last_step_within"),
                          "heldout_free_run": frac("c1_heldout_free_run", "within"),
                          "c2_steps": frac("c2", "steps_within"), "c3_steps": frac("c3", "steps_within")}
        for m, v in (par.get("free_run_other_modes") or {}).items():
            if not isinstance(v, dict):
                continue
            tail["parity"][m + "_free_run"] = r3(v.get("c1"))
            tail["parity"][m + "_heldout_free_run"] = r3(v.get("heldout"))
        try:
            with open(os.path.join(ROOT, "tests", "golden", "noise_floor_bulk.json")) as f:
                tail["parity"]["reference_vs_itself_free_run"] = r3(json.load(f)["summary"]["within"])
        except (OSError, KeyError, ValueError):
            pass
    line["tail"] = tail
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def other_configs(args, rank, world, local_rank, dec, opt, barrier, max_over_ranks, sum_over_ranks):
    """The other BASELINE.json configs under the driver's clock, a few repetitions each (rank 0 returns the keys):
      c1  single KITTI-style car through `reconstruct_object` with host arrays (latency; rank 0 only)
      c3  monocular detections (Redwood hyper-parameters, 48 samples, 20 iterations), 16 per rank
      c4  256^3 mesh grid, slabs over the ranks + overlapped NCCL all-gathers (the one collective of the design)
      c5  1 024 C1-shaped detections split contiguously over the ranks, H2D / D2H included
      parity  statistics of this mode against the reference-made vectors in tests/golden (tools/parity_stats.py)"""
    import torch
    import torch.distributed as dist
    from dspslam import synth
    from dspslam.dist import grid_sdf_sharded, shard_range
    out = {}
    if args.skip_configs:
        return out
    eng = opt.engine
    # ---- the other tensor-core operating points on the same resident batch (rank 0; objects/s per GPU) ----
    if rank == 0:
        prm0 = opt._params()
        batch0 = eng.upload_joint(eng.pack_joint(make_dets(rank)))
        modes = {}
        for m in ("tc_fast", "tc", "tc_exact"):
            for _ in range(2):
                eng.run_joint(batch0, prm0, mode=m)
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
            torch.cuda.synchronize()
            ev[0].record()
            for _ in range(3):
                eng.run_joint(batch0, prm0, mode=m)
            ev[1].record()
            torch.cuda.synchronize()
            modes[m] = OBJ_PER_RANK * 3 / (ev[0].elapsed_time(ev[1]) * 1e-3)
        out["modes_objects_per_s_per_gpu"] = dict(modes, note="C2 batch resident, 3 steps each, no L2 flush; parity of "
                                                  "each mode: profiles/r02_parity_*.jsonl, tests/test_gpu_parity_bulk.py")
        del batch0
    # ---- C1 latency ----
    if rank == 0:
        d = synth.make_kitti_detection(0)
        for _ in range(3):
            o = opt.reconstruct_object(d["t_cam_obj"], d["pts"], d["rays"], d["depth"])
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        reps = 20
        for _ in range(reps):
            o = opt.reconstruct_object(d["t_cam_obj"], d["pts"], d["rays"], d["depth"])
        torch.cuda.synchronize()
        out["c1"] = {"workload": "C1 single KITTI-style car: 250 surface pts + 450 rays x 50 samples, 10 GN iters, "
                                 "reconstruct_object() from host arrays",
                     "c1_ms_per_object": 1e3 * (time.perf_counter() - t0) / reps, "is_good": bool(o.is_good)}
    barrier()
    # ---- C3 mono: render-loss heavy, 20 iterations, 16 detections per rank ----
    from reconstruct.optimizer import Optimizer as _Opt
    opt3 = _Opt(dec, synth.redwood_config(iters=20, num_depth_samples=48), mode=args.mode, device=local_rank)
    dets3 = [synth.make_mono_detection(rank * 16 + k) for k in range(16)]
    for _ in range(2):
        o3 = opt3.reconstruct_objects(dets3)
    barrier()
    t0 = time.perf_counter()
    for _ in range(3):
        o3 = opt3.reconstruct_objects(dets3)
    torch.cuda.synchronize()
    dt3 = max_over_ranks(time.perf_counter() - t0) / 3
    good3 = int(sum_over_ranks(float(sum(1 for o in o3 if o.is_good))))
    out["c3"] = {"workload": "C3 monocular (Redwood hyper-parameters): 16 detections/GPU x 256 surface pts + 1024 rays x 48 "
                             "samples, 20 GN iters, reconstruct_objects() from host arrays",
                 "objects_per_s": 16 * world / dt3, "ms_per_batch": 1e3 * dt3, "good": good3, "of": 16 * world}
    barrier()
    # ---- C4 grid + all-gather ----
    G = 256
    code = np.zeros(64, np.float32)
    code[:3] = [0.2, -0.1, 0.3]
    grid_sdf_sharded(eng, code, G, mode=args.mode)
    barrier()
    tm, tot, comp, expo = {}, [], [], []
    for _ in range(3):
        barrier()
        vol = grid_sdf_sharded(eng, code, G, mode=args.mode, timing=tm)
        tot.append(tm["ms_total"]); comp.append(tm["ms_compute"]); expo.append(tm["ms_allgather_exposed"])
    ms_tot = max_over_ranks(float(np.median(tot)))
    ms_comp = max_over_ranks(float(np.median(comp)))
    ms_expo = max_over_ranks(float(np.median(expo)))
    ms_ag = 0.0
    if world > 1:                                   # the all-gather of the whole volume alone (no compute to hide behind)
        buf = torch.empty(((G ** 3 + world - 1) // world) * world, dtype=torch.float32, device=eng.device)
        per = buf.numel() // world
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        dist.all_gather_into_tensor(buf, buf[rank * per:(rank + 1) * per].clone())
        barrier()
        ev[0].record()
        for _ in range(3):
            dist.all_gather_into_tensor(buf, buf[rank * per:(rank + 1) * per])
        ev[1].record()
        torch.cuda.synchronize()
        ms_ag = max_over_ranks(ev[0].elapsed_time(ev[1]) / 3)
    checksum = float(vol.double().sum().item())
    out["c4"] = {"workload": "C4 mesh grid 256^3 (16.8 M queries): slabs over the ranks, one NCCL all-gather per piece "
                             "overlapping the next piece's kernel",
                 "ms_grid": ms_tot, "ms_slab_compute": ms_comp, "ms_allgather_exposed": ms_expo,
                 "ms_allgather_alone": ms_ag, "ms_allgather_hidden": max(0.0, ms_ag - ms_expo), "pieces": tm.get("pieces"),
                 "mquery_per_s": G ** 3 / ms_tot / 1e3, "tflops": G ** 3 * FLOP_FWD / ms_tot / 1e9,
                 # algorithmic HBM bytes of the grid kernel: 4 B written per voxel (coordinates are generated from the
                 # linear index, nothing is read per row) + the 7.4 MB FP16 weight image once per launch - the kernel is
                 # tensor-bound by three orders of magnitude, the figure is here because north_star asks for it
                 "hbm_gbs_algorithmic": (4.0 * G ** 3 / world + 7.4e6) / (ms_comp * 1e-3) / 1e9, "checksum": checksum}
    del vol
    # ---- C5 sweep ----
    n_total = 1024
    lo, hi = shard_range(n_total, rank, world)
    dets = [synth.make_kitti_detection(k) for k in range(lo, hi)]
    opt5 = opt
    opt5.reconstruct_objects(dets)                  # warm-up pass at the same size (pinned staging, scratch growth)
    barrier()
    t0 = time.perf_counter()
    outs = opt5.reconstruct_objects(dets)
    torch.cuda.synchronize()
    dt = max_over_ranks(time.perf_counter() - t0)
    good = int(sum_over_ranks(float(sum(1 for o in outs if o.is_good))))
    out["c5"] = {"workload": "C5 throughput sweep: 1024 C1-shaped detections split contiguously over the ranks, "
                             "H2D / D2H included, second of two passes", "seconds": dt, "objects_per_s": n_total / dt, "good": good}
    barrier()
    # ---- parity of the benchmarked mode against the reference's vectors ----
    if rank == 0:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import parity_stats
        try:
            out["parity"] = parity_stats.parity(eng, args.mode)
            # the free-running end state of the two modes that reach the reference's own reproducibility floor, for
            # comparison with the benchmarked mode's
            try:
                out["parity"]["free_run_other_modes"] = {
                    m: {k: parity_stats.free_run_stats(eng, m, 128, w)["within"]
                        for k, w in (("c1", "bulk_c1.npz"), ("heldout", "bulk_c1_heldout.npz"))
                        if os.path.isfile(os.path.join(parity_stats.GOLD, w))}
                    for m in ("fp32", "tc_exact") if m != args.mode}
            except Exception as ex:  # noqa: BLE001
                out["parity"]["free_run_other_modes"] = {"error": str(ex)[-200:]}
        except Exception as ex:  # noqa: BLE001
            out["parity"] = {"error": str(ex)[-300:]}
    barrier()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="tc", choices=["tc", "tc_fast", "tc_exact", "fp32"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--skip-configs", action="store_true", help="skip the c1 / c4 / c5 / parity legs")
    ap.add_argument("--ray-phases", type=int, default=0,
                    help="render term: 1 = evaluate every in-ball sample like the reference, 2..4 = early ray "
                         "termination in that many phases (bit-identical results), 0 = library default")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        reference_arm(args, rank, world)
        return
    if world == 1 and args.gpus > 1:
        # launched without torchrun: re-launch under torch.distributed.run
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 2000), os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    ours_arm(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
