"""The reference's OWN scripts run on this package's hot path (BASELINE.json north_star: "orient_simple.py,
orient_pointcloud.py, orient_large.py and reference_orientation.py use it as a drop-in").

Build container only: needs the reference checkout ($DNP_REFERENCE or /root/reference), which does not exist on the
GPU box.  There is no GPU here, so `run(opts)` is followed until its first hot call, which must be THIS package's
(dnp_native raising DipoleError "no CUDA device"), not an ImportError / AttributeError on the way there -- i.e.
the unmodified scripts import, parse their own options, load data/ok.xyz with their own loader, and enter the
library through the reference's call signatures."""
import os
import subprocess
import sys
import textwrap

import pytest

from conftest import PKG

REF = os.environ.get("DNP_REFERENCE", "/root/reference")
pytestmark = pytest.mark.skipif(not os.path.isfile(os.path.join(REF, "orient_simple.py")),
                                reason="reference checkout not present (build container only)")


def _run(body, tmp_path):
    """Each script in a fresh interpreter: the drop-in edits sys.path / sys.modules."""
    code = textwrap.dedent(f"""
        import sys, traceback
        sys.path.insert(0, {PKG!r})
        import torch
        import dnp_dropin, dnp_native
        from pathlib import Path
        REF, OUT = {REF!r}, Path({str(tmp_path)!r})
        def hot_call_reached(fn):
            try:
                fn()
            except dnp_native.DipoleError as e:
                tb = traceback.extract_tb(e.__traceback__)
                files = [f.filename for f in tb]
                assert any(f.startswith(REF) for f in files), files          # the reference's script was running
                assert any(f.endswith("dnp_native.py") for f in files), files
                print("REACHED", [f.name for f in tb if f.filename.startswith(REF)][-1],
                      "->", [f.name for f in tb if {PKG!r} in f.filename][0])
                return
            raise AssertionError("no GPU in this container: the hot call should have raised DipoleError")
    """) + textwrap.dedent(body)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    return r.stdout


@pytest.mark.skipif(__import__("torch").cuda.is_available(), reason="follows the call to the 'no device' error")
def test_orient_simple_reaches_the_library(tmp_path):
    out = _run("""
        dnp_dropin.activate(REF, install_util=False)          # the reference's own text loader and Transform
        mod = dnp_dropin.load("orient_simple")
        import field_utils, options
        assert mod.strongest_field_propagation_points is field_utils.strongest_field_propagation_points
        assert field_utils.__file__.startswith(%r) and mod.util.__file__.startswith(REF)
        opts = options.get_parser().parse_args(["--pc", REF + "/data/ok.xyz", "--export_dir", str(OUT / "o")])
        hot_call_reached(lambda: mod.run(opts))
    """ % PKG, tmp_path)
    assert "REACHED run -> strongest_field_propagation_points" in out


@pytest.mark.skipif(__import__("torch").cuda.is_available(), reason="follows the call to the 'no device' error")
def test_reference_orientation_reaches_the_library(tmp_path):
    out = _run("""
        import argparse
        dnp_dropin.activate(REF, install_util=False)
        mod = dnp_dropin.load("reference_orientation")
        opts = argparse.Namespace(input=Path(REF + "/data/ok.xyz"), reference=Path(REF + "/data/hand.xyz"),
                                  output=OUT / "r.xyz", n=30, estimate_normals=False)
        hot_call_reached(lambda: mod.run(opts))
    """, tmp_path)
    assert "REACHED run -> reference_field" in out


@pytest.mark.skipif(__import__("torch").cuda.is_available(), reason="follows the call to the 'no device' error")
@pytest.mark.parametrize("script", ["orient_large", "orient_pointcloud"])
def test_patch_pipelines_reach_the_library(tmp_path, script):
    """orient_large.py / orient_pointcloud.py import inference_utils -> models.pointcnn -> torch_geometric (not
    installed here: inert placeholders); with `--models` empty their first hot call is the text parse / voxel
    partition this package installs into the reference's util module (its own divide_pc is broken at HEAD)."""
    out = _run(f"""
        dnp_dropin.activate(REF)
        mod = dnp_dropin.load({script!r})
        import options, util, dnp_util
        assert util.divide_pc is dnp_util.divide_pc and util.__file__.startswith(REF)
        opts = options.get_parser().parse_args(["--pc", REF + "/data/ok.xyz", "--export_dir", str(OUT / "l"),
                                                "--number_parts", "10", "--minimum_points_per_patch", "100"])
        hot_call_reached(lambda: mod.run(opts))
    """, tmp_path)
    assert "REACHED run ->" in out


def test_command_line_entry(tmp_path):
    r = subprocess.run([sys.executable, "-m", "dnp_dropin"], capture_output=True, text=True,
                       env=dict(os.environ, PYTHONPATH=PKG, DNP_REFERENCE=REF))
    assert r.returncode != 0 and "usage: python -m dnp_dropin" in r.stderr
