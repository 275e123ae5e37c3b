"""Drop-in for the reference's dipole_api.py: `orient_large(opts)` (:14-87) and
`get_parser()` (:101-132).

`orient_large` is NOT re-implemented here: the pipeline (load -> normals -> partition -> PointCNN votes ->
representative-point propagation -> lattice sign test -> export) is the reference's own control flow, which
north_star leaves unchanged.  The adapter below runs the reference's function object from its checkout with this
package's hot path in front of it (dnp_dropin): its `strongest_field_propagation_reps`, `measure_mean_potential`,
`util.divide_pc`, `util.xyz2tensor` and `util.export_pc` calls land in libdipole.so, everything else (Transform,
orient_center, fix_n_filter, the model loop, MyTimer) executes as written in the reference.

`get_parser` is built from the flag schema recorded from the reference (tests/golden/parser_signatures.json pins it):
same flags, types, defaults and actions; only the fork's hard-coded Windows checkout path is overridable through
$DNP_BASE_PATH.
"""
import argparse
import importlib.util
import os
import sys
from pathlib import Path

import dnp_dropin

_REF_MODULE = "_reference_dipole_api"


def _reference_api():
    """The reference's dipole_api.py as a module object (under a private name: this file shadows `dipole_api`)."""
    mod = sys.modules.get(_REF_MODULE)
    if mod is None:
        ref = dnp_dropin.activate()
        spec = importlib.util.spec_from_file_location(_REF_MODULE, os.path.join(ref, "dipole_api.py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[_REF_MODULE] = mod
        spec.loader.exec_module(mod)
    return mod


def orient_large(opts):
    """dipole_api.py:14-87, executed from the reference checkout ($DNP_REFERENCE or on sys.path) on this package's
    hot path.  Raises FileNotFoundError when no checkout is visible."""
    import field_utils
    field_utils._cuda_device()                    # fail loudly without a GPU before any file is touched
    return _reference_api().orient_large(opts)


# dipole_api.py:93-97: the fork's checkout location, overridable
base_path = os.environ.get("DNP_BASE_PATH", "D:/Documents/zhudoongli/CG/project/NormalEstimation/dipole-normal-prop")
input_pc = Path(base_path + "/data/scene37_0.001_gt68_2.xyz")
models = [Path(base_path + "/pre_trained/hands2.pt"), Path(base_path + "/pre_trained/hands.pt"),
          Path(base_path + "/pre_trained/manmade.pt")]
export_dir = Path(base_path + "/data/output/")

# (flag, type, default, nargs, store_true) -- dipole_api.py:116-128
_FLAGS = (("pc", Path, input_pc, None, False), ("models", Path, models, "+", False),
          ("export_dir", Path, export_dir, None, False), ("number_parts", int, 10, None, False),
          ("minimum_points_per_patch", int, 100, None, False), ("curvature_threshold", float, 0.0, None, False),
          ("iters", int, 5, None, False), ("diffuse", None, True, None, True), ("weighted_prop", None, True, None, True),
          ("estimate_normals", None, True, None, True), ("n", int, 30, None, False))


def get_parser():
    """dipole_api.py:101-132: same flags, types, defaults and actions."""
    parser = argparse.ArgumentParser()
    for name, typ, default, nargs, flag in _FLAGS:
        if flag:
            parser.add_argument("--" + name, default=default, action="store_true")
        elif nargs:
            parser.add_argument("--" + name, type=typ, nargs=nargs, default=default)
        else:
            parser.add_argument("--" + name, type=typ, default=default)
    return parser


if __name__ == '__main__':
    orient_large(get_parser().parse_args())
