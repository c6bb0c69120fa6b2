"""Drop-in for the reference's `extract_map_objects.py:25-63`: same command line, same files written.

    python extract_map_objects.py -c <config.json> -m <map_dir> [-n voxels_dim]

For every object of <map_dir>/MapObjects.txt (id, 3x4 pose, shape code - written by the C++ side,
src/System_util.cc:118-147) writes <map_dir>/objects/<id>.npy (4x4 pose) and <map_dir>/objects/<id>.ply (mesh of the
shape code).  The grid SDF (tcgen05 kernel) and the marching cubes (csrc/mesh.cu) run on the GPU; only vertices and
faces cross to the host."""
import argparse

from dspslam.mesh import extract_map_objects
from reconstruct.optimizer import MeshExtractor
from reconstruct.utils import get_configs, get_decoder


def config_parser():
    parser = argparse.ArgumentParser()
    parser.add_argument('-c', '--config', type=str, required=True, help='path to config file')
    parser.add_argument('-m', '--map_dir', type=str, required=True, help='path to map directory')
    parser.add_argument('-n', '--voxels_dim', type=int, default=128, help='voxels resolution for running marching cube')
    return parser


if __name__ == "__main__":
    args = config_parser().parse_args()
    configs = get_configs(args.config)
    decoder = get_decoder(configs)
    mesh_extractor = MeshExtractor(decoder, configs.optimizer.code_len, args.voxels_dim)
    ids = extract_map_objects(args.map_dir, mesh_extractor)
    print("wrote %d objects to %s/objects" % (len(ids), args.map_dir))
