"""Headless counterpart of the reference's `reconstruct_frame.py` (:26-89): one stereo+LiDAR frame ->
detections (LiDAR crop, mask association, rays) -> joint pose+shape optimisation of every detection -> meshes.

The reference opens an Open3D window (graphics: unavailable here); this script writes `<out>/<k>.ply` (mesh in the
camera frame) and `<out>/objects.txt` instead.  The whole chain runs on the GPU without a host round trip of points or
rays: `FrameWithLiDAR.get_detections` (csrc/frame.cu) -> `Optimizer.reconstruct_frame` -> `MeshExtractor`.

    python reconstruct_frame.py -c config_kitti.json -d <sequence_dir> -i <frame_id> [-o out]   # pre-stored labels
    python reconstruct_frame.py --synthetic 2 [-o out]                                          # synthetic frame
"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from dspslam.mesh import transform_vertices  # noqa: E402
from reconstruct.kitti_sequence import FrameWithLiDAR, KITIISequence  # noqa: E402
from reconstruct.optimizer import MeshExtractor, Optimizer, _engine_for  # noqa: E402
from reconstruct.utils import get_configs, get_decoder, write_mesh_to_ply  # noqa: E402


def config_parser():
    parser = argparse.ArgumentParser()
    parser.add_argument("-c", "--config", type=str, help="path to config file")
    parser.add_argument("-d", "--sequence_dir", type=str, help="path to the KITTI sequence")
    parser.add_argument("-i", "--frame_id", type=int, default=0, help="frame id")
    parser.add_argument("-o", "--out", type=str, default="frame_objects", help="output directory")
    parser.add_argument("--synthetic", type=int, default=None, help="seed of a synthetic frame (no dataset needed)")
    parser.add_argument("--voxels_dim", type=int, default=64)
    return parser


def main(argv=None):
    args = config_parser().parse_args(argv)
    if args.synthetic is not None:
        import types
        from dspslam import synth
        from dspslam.frame_synth import K, T_CAM_VELO, synth_kitti_frame
        configs, decoder = synth.kitti_config(), synth.load_wfit()
        scan, boxes, masks, bbs = synth_kitti_frame(args.synthetic)
        seq = types.SimpleNamespace(configs=configs, K_cam=K, invK_cam=np.linalg.inv(K).astype(np.float32),
                                    T_cam_velo=T_CAM_VELO, engine=_engine_for(decoder))
        frame = FrameWithLiDAR(seq, 0, velo_pts=scan, detections_3d=boxes,
                               det_2d={"pred_masks": masks, "pred_boxes": bbs})
    else:
        configs = get_configs(args.config)
        decoder = get_decoder(configs)
        seq = KITIISequence(args.sequence_dir, configs, decoder=decoder)
        frame = FrameWithLiDAR(seq, args.frame_id)
    optimizer = Optimizer(decoder, configs)
    extractor = MeshExtractor(decoder, configs.optimizer.code_len, args.voxels_dim)

    start = time.time()
    frame.get_detections()
    detections = frame.instances
    results = optimizer.reconstruct_frame(frame.batch) if frame.batch is not None else []
    objects = [(k, det, obj) for k, (det, obj) in enumerate(zip(detections, results)) if obj.is_good]
    print("Reconstructed %d of %d detections, time elapsed: %f seconds" % (len(objects), len(detections),
                                                                         time.time() - start))
    os.makedirs(args.out, exist_ok=True)
    with open(os.path.join(args.out, "objects.txt"), "w") as fh:
        for k, det, obj in objects:
            mesh = extractor.extract_mesh_from_code(obj.code)
            write_mesh_to_ply(transform_vertices(mesh.vertices, obj.t_cam_obj), mesh.faces,
                              os.path.join(args.out, "%d.ply" % k))
            fh.write("%d %d %d %.6f %s\n" % (k, det.num_surface_points, det.rays.shape[0], obj.loss,
                                             " ".join("%.6f" % v for v in obj.t_cam_obj[:3].reshape(-1))))
    return objects


if __name__ == "__main__":
    main()
