"""BASELINE.json configs other than the headline one (bench.py measures configs[1] = C2): one JSON line each.

    python tools/bench_configs.py [c1] [c3] [c4] [c5]                                     (1 GPU)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        tools/bench_configs.py c4 c5                                                      (N GPUs)

  c1  single KITTI-style car through `Optimizer.reconstruct_object` (host arrays in/out): latency per call
  c3  monocular, render-loss heavy: 16 detections x 256 surface pts + 1024 rays x 48 samples, 20 GN iterations
  c4  mesh extraction: 256^3 grid SDF, slabs of the slowest axis per rank + NCCL all-gather, then marching cubes
  c5  throughput sweep: 1024 C1-shaped detections partitioned contiguously across the ranks (no collective);
      objects/s includes packing, H2D of the inputs and D2H of the results
  frame  detection -> optimiser-input builder on a KITTI-sized synthetic frame (120 k LiDAR
      points, 12 boxes, 13 masks of 1242x375): CUDA builder (device time by CUDA events, inputs resident; and the
      public `FrameBuilder.build_kitti` call from host arrays), then frame -> Gauss-Newton without leaving HBM,
      next to the numpy oracle of the same step on the host cores
  torchq  like-for-like GPU data point for the per-query metrics: the decoder queried with STOCK PyTorch FP32 on the
      same B200 the way the reference issues its queries (one shared code expanded and concatenated to (n, 67) rows,
      chunks of 64^3, `no_grad` forward for decode_sdf - loss_utils.py:51-79; forward + one reverse-mode pass with a
      ones cotangent for get_batch_sdf_jacobian - loss_utils.py:82-113; weight-norm hooks active, TF32 off), next to
      this library's kernels on the same rows.  A port of the query pattern, not the reference's files.
  torchgn the whole-object counterpart of torchq: `tools/torch_gn.py` (a stock-PyTorch FP32 port of
      `reconstruct_object`, every tensor on the B200, one detection at a time like the reference) on C1- and C2-shaped
      detections, next to this library on the same detections.
All times are device-synchronised wall times of the public API calls (max over ranks)."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path[:0] = [os.path.join(ROOT, "dsp-slam-rgbd_b200"), ROOT]
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from dspslam import synth  # noqa: E402
from dspslam.dist import grid_sdf_sharded, shard_range  # noqa: E402
from reconstruct.optimizer import MeshExtractor, Optimizer  # noqa: E402

rank = int(os.environ.get("RANK", 0))
world = int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))


def sync():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()


def max_ranks(x):
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def emit(**kw):
    if rank == 0:
        kw["n_gpus"] = world
        print(json.dumps(kw), flush=True)


def timed(fn, reps, warm=2):
    for _ in range(warm):
        fn()
    sync()
    t0 = time.perf_counter()
    for _ in range(reps):
        out = fn()
    torch.cuda.synchronize()
    dt = max_ranks(time.perf_counter() - t0) / reps
    sync()
    return dt, out


dec = synth.load_wfit()
which = [a for a in sys.argv[1:] if a.startswith("c")] or ["c1", "c3", "c4", "c5"]

if "c1" in which and world == 1:
    opt = Optimizer(dec, synth.kitti_config(), mode="tc", device=local)
    d = synth.make_kitti_detection(0)
    dt, out = timed(lambda: opt.reconstruct_object(d["t_cam_obj"], d["pts"], d["rays"], d["depth"]), reps=20, warm=3)
    emit(config="C1 single KITTI-style car: 250 surface pts + 450 rays x 50 samples, 10 GN iters, reconstruct_object()",
         ms_per_object=dt * 1e3, objects_per_s=1.0 / dt, is_good=bool(out.is_good))

if "c3" in which:
    cfg = synth.redwood_config(iters=20, num_depth_samples=48)
    opt = Optimizer(dec, cfg, mode="tc", device=local)
    dets = [synth.make_mono_detection(rank * 16 + k) for k in range(16)]
    dt, outs = timed(lambda: opt.reconstruct_objects(dets), reps=5)
    emit(config="C3 monocular: 16 detections/GPU x 256 surface pts + 1024 rays x 48 samples, 20 GN iters, reconstruct_objects()",
         ms_per_batch=dt * 1e3, objects_per_s=16 * world / dt, good=sum(1 for o in outs if o.is_good))

if "c4" in which:
    G = 256
    ex = MeshExtractor(dec, 64, G, mode="tc", device=local)
    code = np.zeros(64, np.float32)
    code[:3] = [0.2, -0.1, 0.3]
    dt_grid, vol = timed(lambda: grid_sdf_sharded(ex.engine, code, G, coord_mode="reference_shear"), reps=5)
    if rank == 0:
        for _ in range(2):
            v, f = ex.engine.marching_cubes(vol)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        v, f = ex.engine.marching_cubes(vol)
        torch.cuda.synchronize()
        dt_mc = time.perf_counter() - t0
        t0 = time.perf_counter()
        vh, fh = v.cpu().numpy(), f.cpu().numpy()
        dt_d2h = time.perf_counter() - t0
        emit(config="C4 mesh extraction: 256^3 grid SDF (x-slabs + all-gather) + marching cubes on rank 0",
             ms_grid=dt_grid * 1e3, grid_mquery_per_s=G ** 3 / dt_grid / 1e6, grid_tflops=G ** 3 * 3.67104e6 / dt_grid / 1e12,
             ms_marching_cubes=dt_mc * 1e3, ms_mesh_d2h=dt_d2h * 1e3, vertices=int(vh.shape[0]), faces=int(fh.shape[0]),
             mc_gbps=(3 * 4 * G ** 3 + 16 * G ** 3) / dt_mc / 1e9)
    sync()

if "c5" in which:
    n_total = 1024
    lo, hi = shard_range(n_total, rank, world)
    opt = Optimizer(dec, synth.kitti_config(), mode="tc", device=local)
    dets = [synth.make_kitti_detection(k) for k in range(lo, hi)]
    dt, outs = timed(lambda: opt.reconstruct_objects(dets), reps=3, warm=1)
    good = sum(1 for o in outs if o.is_good)
    if world > 1:
        t = torch.tensor([good], dtype=torch.float64, device="cuda")
        dist.all_reduce(t)
        good = int(t.item())
    emit(config="C5 throughput sweep: 1024 C1-shaped detections split contiguously over the GPUs, H2D/D2H included",
         seconds=dt, objects_per_s=n_total / dt, good=good)

if "frame" in sys.argv[1:] and world == 1:
    from dspslam.frame import FrameBuilder
    from dspslam.frame_synth import IMG_H, IMG_W, K, T_CAM_VELO, synth_kitti_frame
    from oracle import frame_oracle as fo          # CPU baseline leg only
    scan, boxes, masks, bbs = synth_kitti_frame(3, n_cars=11, n_ground=110000, pts_per_car=1200)
    opt = Optimizer(dec, synth.kitti_config(), mode="tc", device=local)
    fbuild = FrameBuilder(opt.engine, K, np.linalg.inv(K).astype(np.float32), T_CAM_VELO, IMG_W, IMG_H)
    d_scan = torch.from_numpy(scan).cuda()
    d_masks = torch.from_numpy(masks.view(np.uint8)).cuda()
    d_bbs = torch.from_numpy(bbs).cuda()
    dt_res, fb = timed(lambda: fbuild.build_kitti(d_scan, boxes, d_masks, d_bbs), reps=20, warm=3)
    dt_host, fb = timed(lambda: fbuild.build_kitti(scan, boxes, masks, bbs), reps=10, warm=2)
    # device time of the kernels alone (CUDA events around the C-ABI call; build_kitti ends with one small D2H)
    import ctypes as C
    from dspslam import _native as N
    from dspslam.frame import box_frames
    _, bx, _, _ = box_frames(boxes, T_CAM_VELO)
    base, o = fb.blob.data_ptr(), fb.offs
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]

    def kernels():
        N.check(opt.engine.lib.dspslam_frame_build_kitti(
            opt.engine._h, opt.engine._stream(), C.byref(fbuild.prm), d_scan.data_ptr(), d_scan.shape[0], bx, fb.B,
            d_masks.data_ptr(), d_masks.shape[0], d_bbs.data_ptr(), base + o["pts_off"], base + o["ray_off"],
            base + o["n_fg"], base + o["pts"], base + o["rays"], base + o["depth"], base + o["info"]), "frame")
    for _ in range(3):
        kernels()
    torch.cuda.synchronize()
    ev[0].record()
    for _ in range(50):
        kernels()
    ev[1].record()
    torch.cuda.synchronize()
    dt_k = ev[0].elapsed_time(ev[1]) / 50 * 1e-3
    algo_bytes = scan.nbytes + masks.size + 12 * fb.total_pts + 16 * fb.total_rays    # scan once, masks once, outputs
    dt_gn, outs = timed(lambda: opt.reconstruct_frame(fb), reps=5, warm=2)
    invK = np.linalg.inv(K).astype(np.float32)
    t0 = time.perf_counter()
    for _ in range(3):
        ref = fo.kitti_frame(scan, boxes, masks, bbs, K, invK, T_CAM_VELO, IMG_W, IMG_H)
    dt_cpu = (time.perf_counter() - t0) / 3
    emit(config="frame builder: %d LiDAR points, %d boxes, %d masks %dx%d -> %d surface pts + %d rays"
         % (scan.shape[0], boxes.shape[0], masks.shape[0], IMG_W, IMG_H, fb.total_pts, fb.total_rays),
         us_kernels=dt_k * 1e6, us_build_resident=dt_res * 1e6, us_build_from_host=dt_host * 1e6,
         algorithmic_bytes=int(algo_bytes), kernels_gbps=algo_bytes / dt_k / 1e9,
         ms_frame_gn=dt_gn * 1e3, good=sum(1 for x in outs if x.is_good), with_rays=int(fb.info[:, 4].sum()),
         ms_cpu_numpy_oracle=dt_cpu * 1e3, cpu_cores=os.cpu_count())

if "torchq" in sys.argv[1:] and world == 1:
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    net = synth.load_wfit().cuda().eval()
    eng = Optimizer(dec, synth.kitti_config(), mode="tc", device=local).engine
    code = torch.zeros(64, device="cuda")
    code[:3] = torch.tensor([0.2, -0.1, 0.3])
    n_f, n_j = 128 ** 3, 75776                       # the sizes of tools/prof_kernels.py
    xf = torch.rand((n_f, 3), device="cuda") * 1.6 - 0.8
    xj = torch.rand((n_j, 3), device="cuda") * 1.6 - 0.8

    def torch_forward():
        outs = []
        with torch.no_grad():
            for head in range(0, n_f, 64 ** 3):
                x = xf[head: head + 64 ** 3]
                outs.append(net(torch.cat([code.expand(x.shape[0], -1), x], dim=1)).squeeze(1))
        return torch.cat(outs)

    def torch_jacobian():
        inp = torch.cat([code.expand(n_j, -1), xj], dim=1).unsqueeze(1).requires_grad_(True)
        y = net(inp)
        y.backward(torch.ones_like(y))                # out_dim = 1: the identity cotangent is a vector of ones
        return y.detach(), inp.grad

    def ev_time(fn, reps):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            out = fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps * 1e-3, out

    t_tf, s_t = ev_time(torch_forward, 3)
    t_tj, (y_t, j_t) = ev_time(torch_jacobian, 5)
    cnp = code.cpu().numpy()
    t_of, s_o = ev_time(lambda: eng.sdf(cnp, xf), 5)
    t_oj, (y_o, j_o) = ev_time(lambda: eng.sdf_jac(cnp, xj), 5)
    emit(config="per-query throughput on one B200: stock PyTorch FP32 (cuBLAS SGEMM + autograd) vs this library (tcgen05)",
         torch_forward_mquery_s=n_f / t_tf / 1e6, torch_jacobian_mquery_s=n_j / t_tj / 1e6,
         ours_forward_mquery_s=n_f / t_of / 1e6, ours_jacobian_mquery_s=n_j / t_oj / 1e6,
         forward_speedup=t_tf / t_of, jacobian_speedup=t_tj / t_oj,
         max_abs_sdf_diff=float((s_t - s_o).abs().max()),
         jac_rms_diff=float(((j_t.squeeze(1) - j_o) ** 2).mean().sqrt()))

if "torchgn" in sys.argv[1:] and world == 1:
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import torch_gn
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    net = synth.load_wfit().cuda().eval()
    cfg = synth.kitti_config()
    opt = Optimizer(dec, cfg, mode="tc", device=local)
    for name, M, n_obj in (("C1", 250, 8), ("C2", 2048, 8)):
        dets = [synth.make_kitti_detection(k, M=M, n_bg=200) for k in range(n_obj)]

        def torch_all():
            return [torch_gn.reconstruct_object(net, cfg, d["t_cam_obj"], d["pts"], d["rays"], d["depth"]) for d in dets]

        t_t, res_t = timed(torch_all, reps=2, warm=1)
        t_o, res_o = timed(lambda: opt.reconstruct_objects(dets), reps=5)
        t_o1, _ = timed(lambda: [opt.reconstruct_object(d["t_cam_obj"], d["pts"], d["rays"], d["depth"]) for d in dets], reps=3)
        errs = [synth.pose_errors(a[0].cpu().numpy(), b.t_cam_obj) for a, b in zip(res_t, res_o) if a[2] and b.is_good]
        emit(config="%s-shaped detections (%d surface pts, %d rays x 50 samples, 10 GN iters): stock PyTorch FP32 port "
                    "of reconstruct_object on the same B200, one detection at a time, vs this library" % (name, M, M + 200),
             objects=n_obj, torch_objects_per_s=n_obj / t_t, ours_batched_objects_per_s=n_obj / t_o,
             ours_one_at_a_time_objects_per_s=n_obj / t_o1, speedup_batched=t_t / t_o, speedup_one_at_a_time=t_t / t_o1,
             good_torch=sum(1 for a in res_t if a[2]), good_ours=sum(1 for b in res_o if b.is_good),
             max_translation_diff_m=max(e[0] for e in errs), max_rotation_diff_deg=max(e[1] for e in errs))

if world > 1:
    dist.destroy_process_group()
