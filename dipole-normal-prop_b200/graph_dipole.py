"""The thin per-cloud wrappers of the reference's graph_dipole.py whose signatures are kept
(SURVEY.md section 8a row a10): Transform -> propagate in place -> lattice sign test -> inverse transform.

Only these two functions are provided.  The rest of the reference's graph_dipole.py (patch-graph
construction, the Gurobi MIQP flip solve, the folder batch runner, the import-time mkdir of a Windows
path at graph_dipole.py:9-25) is out of scope and keeps working from the reference checkout, where it
picks up this package's field_utils through `import field_utils`.
"""
import field_utils

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
