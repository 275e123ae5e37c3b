"""Runs the reference's OWN pipeline scripts on the B200 hot path (BASELINE.json north_star: "so that orient_simple.py,
orient_pointcloud.py, orient_large.py and reference_orientation.py use it as a drop-in").

The reference has no plugin API; its boundary is `import field_utils`.  `activate()` arranges the
import path so that this package's field_utils.py (and graph.py / graph_dipole.py / dipole_api.py) shadow the
reference's files of the same name, while everything else -- util, options, inference_utils, models, the scripts
themselves -- stays the reference's, unmodified, from its checkout:

    import dnp_dropin
    dnp_dropin.activate("/path/to/dipole-normal-prop")        # or DNP_REFERENCE=/path/to/...
    orient_simple = dnp_dropin.load("orient_simple")          # the reference's module object
    orient_simple.run(opts)

or from a shell:   python -m dnp_dropin orient_simple --pc data/ok.xyz --export_dir out

What `activate` touches outside the import path, and why (nothing here is arithmetic of the hot path):
  * optional third-party modules the reference imports at module level but this path never calls (open3d: viewers
    and KD-tree normals; pymeshlab; gurobipy: replaced by graph.MIQP here; torch_geometric / torch_cluster: the
    PointCNN estimator) are replaced by inert placeholder modules ONLY when they are not installed, so that
    `import util` / `import inference_utils` succeed.  Calling into a placeholder raises.
  * `torch.symeig` (removed from torch; util.pca_eigen_values and inference_utils.fix_n_filter call it,
    util.py:498, inference_utils.py:57) is mapped to torch.linalg.eigh, which returns the same ascending order.
  * the reference's `util` module gets this package's partition / text-IO functions (dnp_util.install): its own
    `divide_pc` is broken at HEAD (util.py:338-341 returns unmerged nested lists).
"""
import importlib
import importlib.util
import os
import sys
import types

_PKG = os.path.dirname(os.path.abspath(__file__))
_OPTIONAL = ("open3d", "pymeshlab", "gurobipy", "torch_geometric", "torch_cluster", "skimage", "plotly")
_SHADOWED = ("field_utils", "graph", "graph_dipole", "dipole_api")
_state = {"reference": None}


class _Placeholder(types.ModuleType):
    """Stands in for an optional dependency that is not installed: importable, attribute access gives an object
    that raises when it is actually used."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)

        def missing(*a, **k):
            raise ModuleNotFoundError(f"{self.__name__}.{name} was called, but {self.__name__.split('.')[0]} is not installed")
        missing.__name__ = name
        return type(name, (), {"__init__": lambda self_, *a, **k: missing(), "__call__": missing})


class _PlaceholderFinder:
    """sys.meta_path entry serving `root` and every `root.sub.module` of the optional packages that are missing."""

    def __init__(self):
        self.roots = set()

    def find_spec(self, fullname, path=None, target=None):
        if fullname.split(".")[0] not in self.roots:
            return None
        return importlib.util.spec_from_loader(fullname, self, is_package=True)

    def create_module(self, spec):
        mod = _Placeholder(spec.name)
        mod.__path__ = []
        return mod

    def exec_module(self, module):
        pass


_finder = _PlaceholderFinder()


def find_reference(path=None):
    """The reference checkout: `path`, else $DNP_REFERENCE, else the first sys.path entry that holds the
    reference's util.py next to orient_simple.py."""
    for cand in [path, os.environ.get("DNP_REFERENCE")] + list(sys.path):
        if cand and os.path.isfile(os.path.join(cand, "orient_simple.py")) and os.path.isfile(os.path.join(cand, "util.py")) \
                and os.path.abspath(cand) != _PKG:
            return os.path.abspath(cand)
    return None


def activate(reference=None, stub_missing=True, install_util=True):
    """Make the reference's scripts importable with this package's hot path in front.  Returns the checkout path."""
    import torch
    ref = find_reference(reference)
    if ref is None:
        raise FileNotFoundError("reference checkout not found: pass its path or set DNP_REFERENCE")
    for p in (ref, _PKG):                                   # package first, reference second
        while p in sys.path:
            sys.path.remove(p)
    sys.path.insert(0, ref)
    sys.path.insert(0, _PKG)
    for name in _SHADOWED:                                  # a previously imported reference copy must not win
        mod = sys.modules.get(name)
        if mod is not None and os.path.dirname(os.path.abspath(getattr(mod, "__file__", "") or "")) != _PKG:
            del sys.modules[name]
    if stub_missing:
        for name in _OPTIONAL:
            if name in sys.modules or name in _finder.roots:
                continue
            try:
                found = importlib.util.find_spec(name) is not None
            except (ImportError, ValueError):
                found = False
            if not found:
                _finder.roots.add(name)
        if _finder.roots and _finder not in sys.meta_path:
            sys.meta_path.append(_finder)
    try:
        torch.symeig(torch.eye(2))
    except Exception:                                       # removed upstream: same eigenpairs, ascending
        torch.symeig = lambda A, eigenvectors=True, upper=True: torch.linalg.eigh(A)
    cwd = os.getcwd()
    try:
        os.chdir(ref)                                       # graph_dipole.py / util.py open relative paths on import
        util = importlib.import_module("util")
    finally:
        os.chdir(cwd)
    if install_util:
        import dnp_util
        dnp_util.install(util)
    import field_utils
    if getattr(field_utils, "util", None) is not util:     # field_utils was imported before the checkout was visible
        field_utils.util = util
    _state["reference"] = ref
    return ref


def load(script):
    """The reference's module `script` (e.g. "orient_simple"), imported from its checkout with the drop-in active."""
    if _state["reference"] is None:
        activate()
    if script in _SHADOWED:
        raise ValueError(f"{script} is provided by this package; import it directly")
    path = os.path.join(_state["reference"], script + ".py")
    if not os.path.isfile(path):
        raise FileNotFoundError(path)
    spec = importlib.util.spec_from_file_location(script, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[script] = mod
    spec.loader.exec_module(mod)
    return mod


def main(argv=None):
    """python -m dnp_dropin <script> [the script's own flags]: runs the reference's `run(opts)` exactly as its
    `__main__` block does (orient_simple.py:37-43 and the like)."""
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv:
        raise SystemExit("usage: python -m dnp_dropin <orient_simple|orient_pointcloud|orient_large|reference_orientation> ...")
    script, rest = argv[0], argv[1:]
    activate()
    mod = load(script)
    if script == "reference_orientation":                   # its parser lives in the __main__ block (:32-40)
        import argparse
        from pathlib import Path
        parser = argparse.ArgumentParser()
        for flag in ("--input", "--reference", "--output"):
            parser.add_argument(flag, type=Path, required=True)
        parser.add_argument("--n", type=int, default=30)
        parser.add_argument("--estimate_normals", action="store_true")
        return mod.run(parser.parse_args(rest))
    options = importlib.import_module("options")
    opts = options.get_parser().parse_args(rest)
    opts.export_dir.mkdir(exist_ok=True, parents=True)
    options.export_options(opts)
    return mod.run(opts)


if __name__ == "__main__":
    main()
