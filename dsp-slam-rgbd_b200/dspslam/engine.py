"""Batched device engine over libdspslam_sm100.so.

PyTorch is used for device memory, streams and pinned staging only; every kernel that runs
is the library's own.  One `Engine` owns one library handle (packed decoder + scratch) on one GPU.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _native as N
from .packer import check_supported, check_variant, effective_weights

_MODES = {"fp32": N.MODE_FP32, "tc": N.MODE_TC, N.MODE_FP32: N.MODE_FP32, N.MODE_TC: N.MODE_TC,
          "tc_fast": N.MODE_TC_FAST, N.MODE_TC_FAST: N.MODE_TC_FAST,
          "tc_exact": N.MODE_TC_EXACT, N.MODE_TC_EXACT: N.MODE_TC_EXACT,
          "mix_fwd_tc": 2, "mix_jac_tc": 3}     # diagnostic mixes (tools/free_run_stats.py)


_LIVE = []        # weak references to the engines constructed in this process, oldest first


def last_engine(device=None):
    """The most recently constructed live Engine ON `device` (an index; default: the current CUDA device), or None.
    Callers that are handed neither a decoder nor a config - `convert_sdf_voxels_to_mesh`, the sequence classes -
    borrow it.  An engine of another GPU is never returned: a rank helper that holds engines on several devices, or a
    test session with many decoders alive, must not have a frame built or a mesh extracted on the wrong device."""
    if device is None:
        import torch
        device = torch.cuda.current_device() if torch.cuda.is_available() else None
    for ref in reversed(_LIVE):
        eng = ref()
        if eng is not None and getattr(eng, "_h", None) and (device is None or eng.device.index == device):
            return eng
    return None


PINNED_CAP_BYTES = 256 << 20      # bound on the pinned staging an Engine keeps between calls (all threads, all classes)


def _align(n, a=256):
    return (n + a - 1) // a * a


@dataclass
class JointResult:
    t_cam_obj: np.ndarray   # (B, 4, 4) float32
    code: np.ndarray        # (B, L) float32
    status: np.ndarray      # (B,) int32   (0 ok, 1 too few ray samples, 2 NaN loss)
    loss: np.ndarray        # (B,) float32
    dbg: np.ndarray | None = None   # (iters, B, DBG_STRIDE) when requested


class JointBatch:
    """A batch of detections resident in HBM (made by Engine.upload_joint)."""

    def __init__(self, B, L, blob, host, offs, total_pts, total_rays, max_pts, max_rays):
        self.B, self.L = B, L
        self.blob, self.host, self.offs = blob, host, offs
        self.total_pts, self.total_rays = total_pts, total_rays
        self.max_pts, self.max_rays = max_pts, max_rays
        self.h2d_bytes = offs["io_end"]
        self.d2h_bytes = offs["io_end"] - offs["poses"]

    def ptr(self, name):
        return self.blob.data_ptr() + self.offs[name]


class Engine:
    def __init__(self, decoder, device=None, mode="tc"):
        import torch
        self.torch = torch
        if not torch.cuda.is_available():
            raise N.NativeError("dspslam.Engine needs a CUDA device (B200); there is no CPU fallback")
        self.lib = N.lib()
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else
                                   (device if isinstance(device, int) else torch.device(device).index or 0))
        self.mode = _MODES[mode]
        sd = decoder.state_dict() if hasattr(decoder, "state_dict") else decoder
        check_variant(decoder, sd)
        layers = effective_weights(sd)
        self.code_len = check_supported(layers)
        self.in_dim = self.code_len + 3
        desc = N.DecoderDesc(self.code_len, 512, 8, 4)
        Wp = (C.c_void_p * 9)(*[w.ctypes.data for w, _ in layers])
        bp = (C.c_void_p * 9)(*[b.ctypes.data for _, b in layers])
        h = C.c_void_p()
        N.check(self.lib.dspslam_create(C.byref(desc), Wp, bp, self.device.index, C.byref(h)), "dspslam_create")
        self._h = h
        self._keep = layers
        self._pinned = {}
        import weakref
        _LIVE[:] = [r for r in _LIVE if r() is not None] + [weakref.ref(self)]
        if self.mode != N.MODE_FP32 and not self.lib.dspslam_mode_available(h, N.MODE_TC):
            why = (self.lib.dspslam_last_error() or b"").decode()
            self.close()
            raise N.NativeError("mode 'tc' is not available for this decoder: %s.  Construct the Optimizer / Engine with "
                                "mode='fp32' (FFMA kernels) for it." % why)

    # -- lifetime -----------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            self.lib.dspslam_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def launches(self):
        return int(self.lib.dspslam_launch_count(self._h))

    def profile_enable(self, on=True):
        N.check(self.lib.dspslam_profile_enable(self._h, 1 if on else 0), "dspslam_profile_enable")

    def profile_read(self):
        """dict of per-kernel-class device time [ms], rows and launch counts since the last read."""
        p = N.Profile()
        N.check(self.lib.dspslam_profile_read(self._h, C.byref(p)), "dspslam_profile_read")
        return {n: getattr(p, n) for n, _ in N.Profile._fields_}

    # -- helpers ------------------------------------------------------------------------------
    def _staging(self, nbytes):
        """Pinned host staging for one batch.  Buffers are kept per size class (powers of two) and per host thread and
        handed out alternately from a PAIR that is allocated together on the first use of the class, so that no call
        after the first pays for `cudaHostAlloc`; a batch stays valid while the next ONE is packed (callers that keep
        more JointBatch objects alive at once pass cached=False).  The pairs an engine keeps are bounded by
        PINNED_CAP_BYTES: when a new class would take the total over it, the least recently used classes are dropped
        first (their memory goes back once the batches made from them are gone), and a class that does not fit on its
        own is not kept at all - a one-off 1024-detection sweep no longer pins four times its blob for good."""
        import threading
        t = self.torch
        cls = 1 << max(12, int(nbytes - 1).bit_length())
        key = (threading.get_ident(), cls)
        ring = self._pinned.pop(key, None)
        if ring is None:
            if 2 * cls > PINNED_CAP_BYTES:
                return t.empty(cls, dtype=t.uint8, pin_memory=True)[:nbytes]
            kept = sum(2 * k[1] for k in self._pinned)
            for old in list(self._pinned):                       # dicts keep insertion order: oldest use first
                if kept + 2 * cls <= PINNED_CAP_BYTES:
                    break
                kept -= 2 * old[1]
                del self._pinned[old]
            ring = [[t.empty(cls, dtype=t.uint8, pin_memory=True) for _ in range(2)], 0]
        self._pinned[key] = ring                                 # (re-)inserted last = most recently used
        buf = ring[0][ring[1] & 1]
        ring[1] += 1
        return buf[:nbytes]

    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream(self.device).cuda_stream)

    def _dev_f32(self, a):
        t = self.torch
        if isinstance(a, t.Tensor):
            return a.to(device=self.device, dtype=t.float32).contiguous()
        return t.from_numpy(np.ascontiguousarray(np.asarray(a), dtype=np.float32)).to(self.device)

    def _mode(self, mode):
        return self.mode if mode is None else _MODES[mode]

    # -- decoder queries ----------------------------------------------------------------------
    def sdf(self, code, xyz, mode=None):
        """decode_sdf (loss_utils.py:51-79): (n,) device tensor."""
        t = self.torch
        code = self._dev_f32(code)[: self.code_len]
        xyz = self._dev_f32(xyz).reshape(-1, 3)
        out = t.empty(xyz.shape[0], dtype=t.float32, device=self.device)
        N.check(self.lib.dspslam_sdf(self._h, self._stream(), self._mode(mode), code.data_ptr(), xyz.data_ptr(),
                                     xyz.shape[0], out.data_ptr()), "dspslam_sdf")
        return out

    def sdf_jac(self, code, xyz, mode=None):
        """get_batch_sdf_jacobian (loss_utils.py:82-113): (n,), (n, L+3) device tensors."""
        t = self.torch
        code = self._dev_f32(code)[: self.code_len]
        xyz = self._dev_f32(xyz).reshape(-1, 3)
        n = xyz.shape[0]
        sdf = t.empty(n, dtype=t.float32, device=self.device)
        jac = t.empty((n, self.in_dim), dtype=t.float32, device=self.device)
        N.check(self.lib.dspslam_sdf_jac(self._h, self._stream(), self._mode(mode), code.data_ptr(),
                                         xyz.data_ptr(), n, sdf.data_ptr(), jac.data_ptr()), "dspslam_sdf_jac")
        return sdf, jac

    def mean_sdf(self, code, xyz, mode=None):
        t = self.torch
        code = self._dev_f32(code)[: self.code_len]
        xyz = self._dev_f32(xyz).reshape(-1, 3)
        out = t.empty(1, dtype=t.float32, device=self.device)
        N.check(self.lib.dspslam_mean_sdf(self._h, self._stream(), self._mode(mode), code.data_ptr(),
                                          xyz.data_ptr(), xyz.shape[0], out.data_ptr()), "dspslam_mean_sdf")
        return out

    def grid_sdf(self, code, G, i0=0, i1=None, coord_mode="reference_shear", out=None, mode=None):
        """SDF over linear voxel indices [i0, i1) of the G^3 grid (x slowest, z fastest)."""
        t = self.torch
        code = self._dev_f32(code)[: self.code_len]
        i1 = G ** 3 if i1 is None else i1
        if out is None:
            out = t.empty(i1 - i0, dtype=t.float32, device=self.device)
        cm = N.GRID_REFERENCE_SHEAR if coord_mode == "reference_shear" else N.GRID_LATTICE
        N.check(self.lib.dspslam_grid_sdf(self._h, self._stream(), self._mode(mode), code.data_ptr(), G, i0, i1,
                                          cm, out.data_ptr()), "dspslam_grid_sdf")
        return out

    def marching_cubes(self, sdf, level=0.0, spacing=None, origin=-1.0):
        """Iso-surface of a (G,G,G) float32 device tensor: (vertices (nv,3) f32, faces (nf,3) i32) device tensors.
        One counting call sized by a surface-area guess; a second call only if the guess was too small."""
        t = self.torch
        sdf = sdf.to(device=self.device, dtype=t.float32).contiguous()
        G = int(sdf.shape[0])
        if tuple(sdf.shape) != (G, G, G):
            raise ValueError("marching_cubes needs a cubic (G,G,G) grid")
        spacing = 2.0 / (G - 1) if spacing is None else float(spacing)
        counts = t.zeros(3, dtype=t.int32, device=self.device)
        cap_v, cap_t = 8 * G * G + 1024, 16 * G * G + 2048
        while True:
            verts = t.empty((cap_v, 3), dtype=t.float32, device=self.device)
            faces = t.empty((cap_t, 3), dtype=t.int32, device=self.device)
            N.check(self.lib.dspslam_marching_cubes(self._h, self._stream(), sdf.data_ptr(), G, float(level), spacing,
                                                    float(origin), cap_v, cap_t, verts.data_ptr(), faces.data_ptr(),
                                                    counts.data_ptr()), "dspslam_marching_cubes")
            nv, nf, fits = (int(v) for v in counts.tolist())
            if fits:
                return verts[:nv], faces[:nf]
            cap_v, cap_t = nv, nf

    # -- Gauss-Newton --------------------------------------------------------------------------
    def pack_joint(self, dets, pinned=True, cached=False):
        """Concatenate detections into one (pinned) host blob.  `dets`: iterable of dicts/tuples with
        t_cam_obj (4,4), pts (M,3), rays (N,3), depth (n_fg,), optional code.  cached=True takes the blob from the
        engine's ring of pinned staging buffers (the one-shot path of gn_joint: valid until three more batches have
        been packed by this thread); otherwise a fresh pinned allocation the caller owns."""
        t = self.torch
        L = self.code_len
        T, P, R, Dp, Cd = [], [], [], [], []
        for d in dets:
            if isinstance(d, dict):
                T0, pts, rays, depth, code = d["t_cam_obj"], d["pts"], d["rays"], d["depth"], d.get("code")
            else:
                T0, pts, rays, depth = d[:4]
                code = d[4] if len(d) > 4 else None
            T.append(np.asarray(T0, np.float32).reshape(16))
            P.append(np.asarray(pts, np.float32).reshape(-1, 3))
            R.append(np.asarray(rays, np.float32).reshape(-1, 3))
            Dp.append(np.asarray(depth, np.float32).reshape(-1))
            Cd.append(np.zeros(L, np.float32) if code is None else np.asarray(code, np.float32).reshape(-1)[:L])
        B = len(T)
        m = np.array([p.shape[0] for p in P], np.int64)
        n = np.array([r.shape[0] for r in R], np.int64)
        nfg = np.array([d.shape[0] for d in Dp], np.int32)
        if np.any(nfg > n):
            raise ValueError("more depth values than rays")
        pts_off = np.zeros(B + 1, np.int32)
        ray_off = np.zeros(B + 1, np.int32)
        pts_off[1:] = np.cumsum(m)
        ray_off[1:] = np.cumsum(n)
        tp, tr = int(pts_off[-1]), int(ray_off[-1])
        offs, cur = {}, 0
        for name, nbytes in (("pts_off", 4 * (B + 1)), ("ray_off", 4 * (B + 1)), ("n_fg", 4 * B),
                             ("pts", 12 * tp), ("rays", 12 * tr), ("depth", 4 * tr),
                             ("poses0", 64 * B), ("codes0", 4 * L * B),
                             ("poses", 64 * B), ("codes", 4 * L * B), ("status", 4 * B), ("loss", 4 * B)):
            offs[name] = cur
            cur = _align(cur + nbytes)
        offs["io_end"] = cur
        host = self._staging(cur) if (pinned and cached) else t.empty(cur, dtype=t.uint8, pin_memory=pinned)
        hb = host.numpy()

        def put(name, arr):
            b = np.ascontiguousarray(arr).view(np.uint8).reshape(-1)
            hb[offs[name]: offs[name] + b.size] = b

        put("pts_off", pts_off)
        put("ray_off", ray_off)
        put("n_fg", nfg)
        if tp:
            put("pts", np.concatenate(P))
        if tr:
            put("rays", np.concatenate(R))
            depth = np.zeros(tr, np.float32)
            for b in range(B):
                depth[ray_off[b]: ray_off[b] + nfg[b]] = Dp[b]
            put("depth", depth)
        put("poses0", np.stack(T))
        put("codes0", np.stack(Cd))
        return dict(B=B, L=L, host=host, offs=offs, total_pts=tp, total_rays=tr,
                    max_pts=int(m.max()) if B else 0, max_rays=int(n.max()) if B else 0)

    def upload_joint(self, packed):
        blob = packed["host"].to(self.device, non_blocking=True)
        return JointBatch(packed["B"], packed["L"], blob, packed["host"], packed["offs"], packed["total_pts"],
                          packed["total_rays"], packed["max_pts"], packed["max_rays"])

    def run_joint(self, batch, params, mode=None, debug=False):
        """Run the joint GN on a resident batch (asynchronous on the current stream).  Re-runnable:
        every run restarts from the uploaded initial poses/codes.  Returns the debug tensor or None."""
        t = self.torch
        B, L, o = batch.B, batch.L, batch.offs
        io = batch.blob
        io[o["poses"]: o["poses"] + 64 * B].copy_(io[o["poses0"]: o["poses0"] + 64 * B])
        io[o["codes"]: o["codes"] + 4 * L * B].copy_(io[o["codes0"]: o["codes0"] + 4 * L * B])
        prm = N.JointParams(params["k1"], params["k2"], params["k3"], params["k4"], params["b1"], params["b2"],
                            params["lr"], params["s_damp"], params["cut_off"], params["num_iterations"],
                            params["num_depth_samples"], int(params.get("ray_phases", 0)),
                            float(params.get("refine_window", 0.0)))
        dbg = None
        if debug:
            dbg = t.zeros((max(1, params["num_iterations"]), B, N.DBG_STRIDE), dtype=t.float32, device=self.device)
        N.check(self.lib.dspslam_gn_joint(
            self._h, self._stream(), self._mode(mode), C.byref(prm), B, batch.ptr("pts_off"), batch.ptr("ray_off"),
            batch.ptr("n_fg"), batch.ptr("pts"), batch.ptr("rays"), batch.ptr("depth"), batch.total_pts,
            batch.total_rays, batch.max_pts, batch.max_rays, batch.ptr("poses"), batch.ptr("codes"),
            batch.ptr("status"), batch.ptr("loss"), dbg.data_ptr() if dbg is not None else None), "dspslam_gn_joint")
        return dbg

    def download_joint(self, batch, dbg=None):
        t = self.torch
        o, B, L = batch.offs, batch.B, batch.L
        lo, hi = o["poses"], o["io_end"]
        batch.host[lo:hi].copy_(batch.blob[lo:hi], non_blocking=True)
        t.cuda.current_stream(self.device).synchronize()
        hb = batch.host.numpy()

        def get(name, dtype, count):
            return hb[o[name]: o[name] + count * np.dtype(dtype).itemsize].view(dtype).copy()

        return JointResult(get("poses", np.float32, 16 * B).reshape(B, 4, 4), get("codes", np.float32, L * B).reshape(B, L),
                           get("status", np.int32, B), get("loss", np.float32, B),
                           None if dbg is None else dbg.cpu().numpy())

    def gn_joint(self, dets, params, mode=None, debug=False):
        """Host arrays in, host arrays out (the call `reconstruct_object(s)` makes)."""
        batch = self.upload_joint(self.pack_joint(dets, cached=True))
        dbg = self.run_joint(batch, params, mode=mode, debug=debug)
        return self.download_joint(batch, dbg)

    def gn_pose_only_resident(self, batch, t_co_se3, scales, codes, num_iterations, mode=None):
        """Pose-only GN on the surface points of a resident batch (JointBatch / FrameBatch: `pts_off`, `pts` already in
        HBM): only the (B,4,4) SE(3) poses, B scales and B codes go up, (B,4,4) comes back."""
        t = self.torch
        B, L = batch.B, self.code_len
        T = np.ascontiguousarray(np.asarray(t_co_se3, np.float32).reshape(B, 16))
        sc = np.ascontiguousarray(np.asarray(scales, np.float32).reshape(B))
        Cd = np.ascontiguousarray(np.asarray(codes, np.float32).reshape(B, -1)[:, :L])
        if Cd.shape[1] < L:
            Cd = np.concatenate([Cd, np.zeros((B, L - Cd.shape[1]), np.float32)], 1)
        d_T, d_sc, d_code = (t.from_numpy(a).to(self.device) for a in (T, sc, Cd))
        N.check(self.lib.dspslam_gn_pose_only(
            self._h, self._stream(), self._mode(mode), num_iterations, B, batch.ptr("pts_off"), batch.ptr("pts"),
            batch.total_pts, batch.max_pts, d_sc.data_ptr(), d_code.data_ptr(), d_T.data_ptr(), None),
            "dspslam_gn_pose_only")
        return d_T.cpu().numpy().reshape(B, 4, 4)

    def gn_pose_only(self, items, num_iterations, mode=None, debug=False):
        """items: iterable of (t_co_se3 (4,4), scale, pts (M,3), code).  Returns (B,4,4) float32 [, dbg]."""
        t = self.torch
        L = self.code_len
        items = list(items)
        T = np.stack([np.asarray(i[0], np.float32).reshape(16) for i in items])
        sc = np.array([float(i[1]) for i in items], np.float32)
        P = [np.asarray(i[2], np.float32).reshape(-1, 3) for i in items]
        Cd = np.stack([np.asarray(i[3], np.float32).reshape(-1)[:L] for i in items])
        B = T.shape[0]
        pts_off = np.zeros(B + 1, np.int32)
        pts_off[1:] = np.cumsum([p.shape[0] for p in P])
        # one pinned staging blob, one H2D copy (five pageable copies used to cost more than the five GN iterations)
        pts_all = np.concatenate(P)
        offs, cur = {}, 0
        for name, arr in (("off", pts_off), ("pts", pts_all), ("T", T), ("sc", sc), ("code", np.ascontiguousarray(Cd))):
            offs[name] = cur
            cur = _align(cur + arr.nbytes)
        host = self._staging(cur)
        hb = host.numpy()
        for name, arr in (("off", pts_off), ("pts", pts_all), ("T", T), ("sc", sc), ("code", np.ascontiguousarray(Cd))):
            b8 = np.ascontiguousarray(arr).view(np.uint8).reshape(-1)
            hb[offs[name]: offs[name] + b8.size] = b8
        blob = host.to(self.device, non_blocking=True)
        base = blob.data_ptr()
        dbg = t.zeros((max(1, num_iterations), B, 64), dtype=t.float32, device=self.device) if debug else None
        N.check(self.lib.dspslam_gn_pose_only(
            self._h, self._stream(), self._mode(mode), num_iterations, B, base + offs["off"], base + offs["pts"],
            int(pts_off[-1]), int(max(p.shape[0] for p in P)), base + offs["sc"], base + offs["code"], base + offs["T"],
            dbg.data_ptr() if dbg is not None else None), "dspslam_gn_pose_only")
        lo = offs["T"]
        host[lo: lo + 64 * B].copy_(blob[lo: lo + 64 * B], non_blocking=True)
        t.cuda.current_stream(self.device).synchronize()
        out = hb[lo: lo + 64 * B].view(np.float32).reshape(B, 4, 4).copy()
        return (out, dbg.cpu().numpy()) if debug else out
