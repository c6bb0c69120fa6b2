"""Multi-GPU plumbing (one process per GPU, torch.distributed): SURVEY.md 8(e).

* Detections are independent (`src/LocalMapping_util.cc:165-206` loops over them sequentially), so they are
  split contiguously over the ranks and optimised with NO data-path collective; only the results
  (4x4 pose + code + status + loss per object) are gathered on the host side.
* The mesh grid (`reconstruct/optimizer.py:224-228`) is split into contiguous slabs of linear voxel indices -
  slabs of the slowest axis x, the contiguous equivalent of the z-slabs the reference layout would make strided
  (`reconstruct/utils.py:106`: z is the fastest axis) - evaluated with `dspslam_grid_sdf(i0, i1)` on each rank
  and gathered with one all-gather (NCCL over NVLink on the B200 box; gloo in the CPU tests).
"""
from __future__ import annotations

import numpy as np


def shard_range(n, rank, world):
    """Contiguous [lo, hi) of `n` items owned by `rank`: sizes differ by at most one, earlier ranks larger."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def slab_range(total, rank, world):
    """Equal-size slabs (all-gather needs equal counts): [lo, hi) clipped to `total`, and the padded slab size."""
    per = (total + world - 1) // world
    lo = min(rank * per, total)
    return lo, min(lo + per, total), per


def _world(group):
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def grid_sdf_sharded(engine, code, G, coord_mode="reference_shear", mode=None, group=None):
    """(G, G, G) SDF volume on every rank; each rank evaluates one slab, one all-gather assembles it.
    `engine` needs `.grid_sdf(code, G, i0, i1, coord_mode=, out=, mode=)` and `.torch`/`.device` (dspslam.Engine)."""
    import torch
    import torch.distributed as dist
    rank, world = _world(group)
    total = G ** 3
    lo, hi, per = slab_range(total, rank, world)
    full = torch.empty(per * world, dtype=torch.float32, device=engine.device)
    mine = full[rank * per: rank * per + per]
    if hi > lo:
        engine.grid_sdf(code, G, i0=lo, i1=hi, coord_mode=coord_mode, out=mine[: hi - lo], mode=mode)
    if world > 1:
        dist.all_gather_into_tensor(full, mine.clone() if full.device.type == "cpu" else mine, group=group)
    return full[:total].view(G, G, G)


def reconstruct_sharded(optimizer, detections, group=None):
    """`Optimizer.reconstruct_objects` over this rank's contiguous shard of `detections`; returns the full result
    list on every rank (host-side gather of the small per-object results; no data-path collective)."""
    import torch.distributed as dist
    rank, world = _world(group)
    lo, hi = shard_range(len(detections), rank, world)
    mine = optimizer.reconstruct_objects(detections[lo:hi]) if hi > lo else []
    if world == 1:
        return mine
    packed = [(bool(o.is_good), None if o.t_cam_obj is None else np.asarray(o.t_cam_obj),
               None if o.code is None else np.asarray(o.code), float(o.loss)) for o in mine]
    parts = [None] * world
    dist.all_gather_object(parts, packed, group=group)
    from reconstruct.utils import ForceKeyErrorDict
    return [ForceKeyErrorDict(is_good=g, t_cam_obj=T, code=c, loss=l) for part in parts for (g, T, c, l) in part]
