"""The drop-in mechanism (dnp_dropin) end to end ON THE GPU.

The reference checkout does not travel to the GPU box, so this test writes a miniature stand-in checkout into a
temporary directory -- `util.py`, `options.py` and a pipeline script with the module layout the reference's scripts
have (`from options import get_parser`, `from field_utils import *`, `util.xyz2tensor`, `util.Transform.trans`, the
greedy propagation, the lattice test, `util.export_pc`; orient_simple.py:1-34) -- and runs it exactly the way a user of
the reference would: `python -m dnp_dropin orient_simple --pc cloud.xyz --export_dir out` with DNP_REFERENCE pointing
at the checkout.  The result file is compared with the CPU oracle's decisions for the same cloud.
(tests/test_dropin.py runs the reference's REAL scripts through the same mechanism in the build container, where
there is no GPU: there the call is followed up to the library's "no device" error.)
"""
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "dipole-normal-prop_b200")
sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu

STANDIN_UTIL = '''
    import contextlib
    import torch

    def xyz2tensor(txt, append_normals=True):
        raise AssertionError("the stand-in loader must have been replaced by the GPU text parser")

    def export_pc(pc, dest):
        raise AssertionError("the stand-in writer must have been replaced by the GPU text formatter")

    class Transform:
        """centre + isotropic scale of the positions into the unit box; normals untouched"""
        def __init__(self, centre, scale):
            self.centre, self.scale = centre, scale

        @classmethod
        def trans(cls, pc):
            centre = pc[:, :3].mean(dim=0)
            scale = (pc[:, :3] - centre).abs().max() * 2.5
            out = pc.clone()
            out[:, :3] = (pc[:, :3] - centre) / scale
            return out, cls(centre, scale)

        def inverse(self, pc):
            out = pc.clone()
            out[:, :3] = pc[:, :3] * self.scale + self.centre
            return out

    def timer_factory():
        class Timer(contextlib.ContextDecorator):
            def __init__(self, label, count=True):
                self.label = label
            def __enter__(self):
                return self
            def __exit__(self, *exc):
                return False
            @staticmethod
            def print_total_time():
                pass
        return Timer
'''

STANDIN_OPTIONS = '''
    import argparse
    from pathlib import Path

    def get_parser():
        p = argparse.ArgumentParser()
        p.add_argument("--pc", type=Path)
        p.add_argument("--export_dir", type=Path)
        p.add_argument("--diffuse", action="store_true")
        return p

    def export_options(opts):
        (opts.export_dir / "options.txt").write_text(repr(vars(opts)))
'''

STANDIN_SCRIPT = '''
    from options import get_parser
    from field_utils import *

    def run(opts):
        device = torch.device("cuda", torch.cuda.current_device())
        Timer = util.timer_factory()
        with Timer("load"):
            cloud = util.xyz2tensor(open(opts.pc).read()).to(device)
        cloud, transform = util.Transform.trans(cloud)
        with Timer("propagate"):
            strongest_field_propagation_points(cloud, diffuse=opts.diffuse, starting_point=0)
        with Timer("global sign"):
            if measure_mean_potential(cloud) < 0:
                cloud[:, 3:] *= -1
        util.export_pc(transform.inverse(cloud).transpose(0, 1), opts.export_dir / "final_result.xyz")
'''


def _cloud(n, seed):
    """an ellipsoid surface with outward normals, half of them flipped"""
    g = torch.Generator().manual_seed(seed)
    d = torch.nn.functional.normalize(torch.randn(n, 3, generator=g), dim=1)
    axes = torch.tensor([1.0, 0.7, 0.5])
    pos = d * axes
    nrm = torch.nn.functional.normalize(d / axes, dim=1)
    flip = (torch.rand(n, generator=g) < 0.5).float() * 2 - 1
    return torch.cat([pos, nrm * flip[:, None]], dim=1)


@pytest.mark.parametrize("diffuse", [False, True])
def test_pipeline_script_runs_through_the_dropin(tmp_path, diffuse):
    checkout = tmp_path / "checkout"
    checkout.mkdir()
    (checkout / "util.py").write_text(textwrap.dedent(STANDIN_UTIL))
    (checkout / "options.py").write_text(textwrap.dedent(STANDIN_OPTIONS))
    (checkout / "orient_simple.py").write_text(textwrap.dedent(STANDIN_SCRIPT))
    pc = _cloud(600, 5)
    cloud_file = tmp_path / "cloud.xyz"
    np.savetxt(cloud_file, pc.numpy(), fmt="%.9g")
    pc = torch.from_numpy(np.loadtxt(cloud_file, dtype=np.float64).astype(np.float32))    # what the parser reads

    env = dict(os.environ, DNP_REFERENCE=str(checkout), PYTHONPATH=PKG)
    cmd = [sys.executable, "-m", "dnp_dropin", "orient_simple", "--pc", str(cloud_file), "--export_dir",
           str(tmp_path / "out")] + (["--diffuse"] if diffuse else [])
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=str(tmp_path))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert (tmp_path / "out" / "options.txt").is_file()
    got = np.loadtxt(tmp_path / "out" / "final_result.xyz", dtype=np.float64).astype(np.float32)
    assert got.shape == (600, 6)

    # the same pipeline on the CPU oracle
    from oracle import torch_drivers
    cpu = pc.clone()
    centre = cpu[:, :3].mean(dim=0)
    scale = (cpu[:, :3] - centre).abs().max() * 2.5
    cpu[:, :3] = (cpu[:, :3] - centre) / scale
    torch_drivers.propagate_points(cpu, diffuse=diffuse, starting_point=0)
    if torch_drivers.lattice_mean_potential(cpu) < 0:
        cpu[:, 3:] *= -1
    # decisions are signs: every normal is the input normal or its exact negation, the same one as the oracle's
    assert np.array_equal(got[:, 3:], cpu[:, 3:].numpy())
    # and the result is the outward orientation of the ellipsoid (or all inward before the lattice test fixed it)
    outward = (torch.from_numpy(got[:, 3:]) * torch.nn.functional.normalize(pc[:, :3] / torch.tensor([1.0, 0.49, 0.25]), dim=1)).sum(1)
    assert (outward > 0).float().mean() > 0.97
    np.testing.assert_allclose(got[:, :3], pc[:, :3].numpy(), rtol=0, atol=2e-6)
