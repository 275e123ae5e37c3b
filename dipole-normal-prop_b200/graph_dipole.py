"""The thin per-cloud wrappers of the reference's graph_dipole.py whose signatures are kept
(SURVEY.md section 8a row a10): Transform -> propagate in place -> lattice sign test -> inverse transform.

Provided: the per-cloud wrappers (:27-42) and graph_dipole_server_api (:87-117), the grid-partition ->
per-patch orientation -> flip scores -> flip solve pipeline the fork's server calls.  Not provided: the
file/folder batch runners around them (open3d IO, ground-truth metrics, viewers) and the import-time
mkdir of a Windows path (:9-25).
"""
import numpy as np
import torch

import dnp_util
import field_utils
from graph import *        # noqa: F401,F403  (the reference does the same, graph_dipole.py:6)
from graph import MIQP

util = field_utils.util


def propagate_points_file(input_pc, propagate_func, *args, **kwargs):
    """graph_dipole.py:27-33."""
    input_pc, transform = util.Transform.trans(input_pc)
    propagate_func(input_pc, *args, **kwargs)
    if field_utils.measure_mean_potential(input_pc) < 0:
        input_pc[:, 3:] *= -1
    input_pc = transform.inverse(input_pc)
    return input_pc


def st_propagation_points_file(input_pc, verbose=True):
    """graph_dipole.py:35-36."""
    return propagate_points_file(input_pc, field_utils.strongest_field_propagation_points, diffuse=True,
                                 starting_point=0, verbose=verbose)


def st_propagation_points_batched(pc_in, indices):
    """st_propagation_points_file on every patch of `indices` at once, in place on pc_in [N, 6] float32 (CUDA):
    the per-patch Transform (util.py:577-609) as segment reductions, ONE launch with a thread-block cluster per
    patch for the greedy propagation (dnp_points_propagate_batched_f32), ONE launch for the per-patch lattice
    sign test (dnp_patch_lattice_sign_f32), the inverse Transform, one scatter.  The reference runs this stage
    as one Python thread per patch (util.py:185-196)."""
    import dnp_native as _dn
    dev = pc_in.device
    sizes = [int(len(i)) for i in indices]
    if not sizes:
        return pc_in
    counts = torch.tensor(sizes, dtype=torch.int64, device=dev)
    offsets = torch.zeros(len(sizes) + 1, dtype=torch.int64, device=dev)
    offsets[1:] = torch.cumsum(counts, 0)
    perm = torch.cat([i.to(dev) for i in indices])
    seg = torch.repeat_interleave(torch.arange(len(sizes), device=dev), counts)
    work = pc_in[perm].contiguous()
    xyz = work[:, :3].contiguous()
    center = torch.segment_reduce(xyz, "mean", lengths=counts, axis=0)
    extent = torch.segment_reduce(xyz, "max", lengths=counts, axis=0) - torch.segment_reduce(xyz, "min", lengths=counts, axis=0)
    scale = extent.max(dim=1).values
    work[:, :3] = (xyz - center[seg]) / scale[seg, None]
    _dn.points_propagate_batched(work, offsets, max(sizes), diffuse=True, eps=field_utils._EPS_POINTS)
    _dn.patch_lattice_sign(work, offsets, 10, flip=True)
    work[:, :3] = work[:, :3] * scale[seg, None] + center[seg]
    pc_in[perm] = work
    return pc_in


st_propagation_points_file.batched = st_propagation_points_batched


def xie_tree_propagation_points_file(input_pc, eps=1e-2, verbose=True, times=1, use_pw=False):
    """graph_dipole.py:41-42."""
    return propagate_points_file(input_pc, field_utils.xie_propagation_points_onbfstree, eps=eps, diffuse=True,
                                 starting_point=0, verbose=verbose, times=times, use_pw=use_pw, knn_mask=-1)


def graph_dipole_server_api(xyz_data, config):
    """graph_dipole.py:87-117 for config['divide_method'] == 'grid_partition': normals -> Transform -> patch graph
    (per-patch greedy orientation, one batched flip-score launch) -> flip solve -> flipped patches; returns the
    [N, 6] numpy array.  Divergences: the cloud is cast to float32 (the reference keeps the wire's float64 and
    then fails in torch.searchsorted on mixed dtypes, util.py:93); 'ncut_partition' needs the fork's external
    toolbox.pointcloud_segmentation package and is not provided."""
    field_utils._cuda_device()
    device = torch.device("cuda", torch.cuda.current_device())
    input_pc = dnp_util.npxyz2tensor(np.asarray(xyz_data)).to(device)
    try:            # the reference's estimator (open3d KD-tree PCA, util.py:551-567) when it is really there
        import open3d
        input_pc = util.estimate_normals(input_pc, max_nn=config['max_nn']).float()
    except (ImportError, AttributeError, ModuleNotFoundError):
        input_pc = dnp_util.knn_pca_normals(input_pc.float(), k=config['max_nn'])
    input_pc, transform = util.Transform.trans(input_pc)
    if config['divide_method'] == 'grid_partition':
        G, index = dnp_util.divide_pc_to_graph(input_pc, n_part=config["n_part"], min_patch=config["min_patch"],
                                               edge_calculator=field_utils.field_edge_calculator,
                                               point_estimator=st_propagation_points_file)
    elif config['divide_method'] == 'ncut_partition':
        raise NotImplementedError("ncut_partition needs toolbox.pointcloud_segmentation (not part of the reference checkout)")
    else:
        print("Error: no such divide method")
        return
    input_pc = transform.inverse(input_pc)
    A, B = G.to_matrix()
    flip = MIQP(A, B)
    for i in range(len(flip)):
        if flip[i] == 1:
            input_pc[index[i], 3:] *= -1
    return input_pc.cpu().numpy()
