"""Drop-in for the reference's dipole_api.py: `orient_large(opts)` (:14-87) and
`get_parser()` (:101-132), the representative-point pipeline of orient_large.py behind one function.

Every stage on the dipole path runs through this package: text parse (dnp_xyz_parse_f32), voxel partition
(dnp_voxel_partition_f32 + merge), representative-point propagation (strongest_field_propagation_reps), the lattice
sign test and the text export (dnp_xyz_format_f32).  The PointCNN stage is the reference's own: its loader
(inference_utils.load_model_from_file, which needs torch_geometric) is imported lazily from the reference checkout
when `opts.models` is non-empty; with no models the stage is skipped, everything else is unchanged.
"""
import argparse
import os
from pathlib import Path

import torch

import dnp_util
import field_utils
from field_utils import *        # noqa: F401,F403  (dipole_api.py:9)

util = field_utils.util
torch.manual_seed(1)             # dipole_api.py:8

estimate_normals = True


def voting_policy(probs):
    """inference_utils.py:46-48."""
    probs = torch.stack(probs, dim=0).mean(dim=0)
    return probs < 0.5, probs


def fix_n_filter(input_pc, patch_indices, threshold):
    """inference_utils.py:51-73: keep the patches whose flatness ratio e0 / (e1 + e2 / 2) exceeds `threshold`;
    the others are oriented along their PCA normal in place.  Returns [(i, patch)]."""
    kept = []
    for i, patch in enumerate(patch_indices):
        x = input_pc[patch]
        temp = x[:, :3] - x.mean(dim=0)[None, :3]
        cov = (temp.transpose(0, 1) @ temp) / x.shape[0]
        e, v = torch.linalg.eigh(cov)
        if (e[0] / ((e[1] + e[2] / 2))).item() > threshold:
            kept.append((i, patch))
        else:
            sign = ((input_pc[patch, 3:] * v[:, 0][None, :]).sum(dim=-1) > 0).float() * 2 - 1
            input_pc[patch, 3:] = input_pc[patch, 3:] * sign[:, None]
    return kept


def _load_models(paths, device):
    if not paths:
        return []
    from inference_utils import load_model_from_file      # the reference's PointCNN loader (torch_geometric)
    return [load_model_from_file(p, device) for p in paths]


def orient_large(opts):
    """dipole_api.py:14-87."""
    field_utils._cuda_device()
    export_path: Path = opts.export_dir
    export_path.mkdir(exist_ok=True)

    max_patch_size = 500
    device = torch.device("cuda", torch.cuda.current_device())
    MyTimer = dnp_util.timer_factory()
    with MyTimer('load pc', count=False):
        input_pc = dnp_util.xyz2tensor_cuda(open(opts.pc, 'r').read(), append_normals=False, device=device)
    print(input_pc.shape)

    input_pc, transform = util.Transform.trans(input_pc)

    if opts.estimate_normals:
        with MyTimer('estimating normals'):
            input_pc = getattr(util, "estimate_normals", dnp_util.estimate_normals)(input_pc, max_nn=opts.n)

    softmax = torch.nn.Softmax(dim=-1)
    models = _load_models(list(opts.models), device)

    with MyTimer('divide patches'):
        # upstream composition: util.divide_pc as shipped returns unmerged nested lists
        patch_indices = dnp_util.merge_nodes(input_pc, *dnp_util._divide_pc(input_pc[:, :3], opts.number_parts,
                                                                             min_patch=opts.minimum_points_per_patch),
                                             opts.minimum_points_per_patch)[0]
        all_patches_indices = [x.clone() for x in patch_indices]

    with MyTimer('filter patches'):
        patch_indices = fix_n_filter(input_pc, patch_indices, opts.curvature_threshold)

    num_patches = len(patch_indices)
    num_all_patches = len(all_patches_indices)
    print(f'number of patches {num_patches}/{num_all_patches}')

    with MyTimer('orient center'):
        for i, p in patch_indices:
            input_pc[p] = dnp_util.orient_center(input_pc[p])

    with MyTimer('find reps'):
        represent = []
        for p in all_patches_indices:
            permutation = torch.randperm(p.shape[0]).to(p.device)
            represent.append((p[permutation[:max_patch_size]], p[permutation[max_patch_size:]]))

    pc_probs = torch.ones_like(input_pc[:, 0])

    with MyTimer('network orientation'):
        if models:
            for i, _ in patch_indices:
                with torch.no_grad():
                    current_reps, non_reps_points = represent[i]
                    data = input_pc[current_reps].to(device)
                    for _ in range(opts.iters):
                        votes = [model(data.clone()) for model in models]
                        vote_probabilities = [softmax(scores)[:, 1] for scores in votes]
                        flip, probs = voting_policy(vote_probabilities)
                        pc_probs[current_reps] = probs
                        input_pc[current_reps[flip], 3:] *= -1

    [model.to('cpu') for model in models]
    with MyTimer('propagating field'):
        strongest_field_propagation_reps(input_pc, represent, diffuse=True)      # noqa: F405

    with MyTimer('fix global orientation'):
        if field_utils.measure_mean_potential(input_pc) < 0:
            input_pc[:, 3:] *= -1

    with MyTimer('exporting result', count=False):
        dnp_util.export_pc(transform.inverse(input_pc).transpose(0, 1), export_path / 'final_result.xyz')

    MyTimer.print_total_time()


# dipole_api.py:93-97: the fork's checkout location, overridable
base_path = os.environ.get("DNP_BASE_PATH", "D:/Documents/zhudoongli/CG/project/NormalEstimation/dipole-normal-prop")
input_pc = Path(base_path + "/data/scene37_0.001_gt68_2.xyz")
models = [Path(base_path + "/pre_trained/hands2.pt"), Path(base_path + "/pre_trained/hands.pt"),
          Path(base_path + "/pre_trained/manmade.pt")]
export_dir = Path(base_path + "/data/output/")


def get_parser():
    """dipole_api.py:101-132: same flags, types and defaults."""
    parser = argparse.ArgumentParser()
    parser.add_argument('--pc', type=Path, default=input_pc, help='path to the point cloud')
    parser.add_argument('--models', type=Path, nargs='+', default=models, help='path to the models')
    parser.add_argument('--export_dir', type=Path, default=export_dir, help='path to the output directory')
    parser.add_argument('--number_parts', type=int, default=10, help='number of parts to divide the point cloud')
    parser.add_argument('--minimum_points_per_patch', type=int, default=100, help='minimum number of points per patch')
    parser.add_argument('--curvature_threshold', type=float, default=0.0, help='curvature threshold')
    parser.add_argument('--iters', type=int, default=5, help='number of iterations')
    parser.add_argument('--diffuse', default=True, action='store_true', help='diffuse')
    parser.add_argument('--weighted_prop', default=True, action='store_true', help='weighted propagation')
    parser.add_argument('--estimate_normals', default=True, action='store_true', help='estimate normals')
    parser.add_argument('--n', type=int, default=30, help='number of nearest neighbors for normal estimation')
    return parser


if __name__ == '__main__':
    orient_large(get_parser().parse_args())
