"""dipole_api mirror: parser parity on CPU, orient_large end to end on the GPU."""
import json
import os

import numpy as np
import pytest
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _describe(parser):
    rows = []
    for a in parser._actions:
        if a.dest == "help":
            continue
        rows.append({"flags": a.option_strings, "dest": a.dest, "type": getattr(a.type, "__name__", None),
                     "default": repr(a.default), "nargs": a.nargs, "action": type(a).__name__, "required": a.required})
    return rows


def test_get_parser_matches_the_reference():
    import dipole_api
    want = json.load(open(os.path.join(GOLDEN, "parser_signatures.json")))["dipole_api.get_parser"]
    assert _describe(dipole_api.get_parser()) == want


@pytest.mark.gpu
def test_orient_large_end_to_end(dev, tmp_path):
    """Text in -> partition -> representative-point propagation -> lattice sign test -> text out, every stage on
    the GPU path: a sphere with scrambled normal signs must come back uniformly outward (positive mean lattice
    potential = outward, field_utils.py:7-9), positions unchanged."""
    import dipole_api
    import dnp_util
    g = torch.Generator().manual_seed(21)
    d = torch.randn(20000, 3, generator=g)
    d = d / d.norm(dim=-1, keepdim=True)
    pc = torch.cat([d * 1.7 + torch.tensor([0.3, -0.2, 0.1]), d], 1).float()
    # coherent scramble: flip whole octants, as a network that votes per patch would leave them
    octant = (d > 0).long() @ torch.tensor([1, 2, 4])
    pc[(octant % 3 == 0), 3:] *= -1
    src = tmp_path / "sphere.xyz"
    dnp_util.export_pc(pc.transpose(0, 1), src)
    opts = dipole_api.get_parser().parse_args(["--pc", str(src), "--export_dir", str(tmp_path / "out"),
                                               "--number_parts", "8", "--minimum_points_per_patch", "60"])
    opts.models = []                      # no PointCNN weights in this test: the network stage is skipped
    opts.estimate_normals = False         # the file carries normals
    dipole_api.orient_large(opts)
    out = dnp_util.xyz2tensor(open(tmp_path / "out" / "final_result.xyz").read())
    assert out.shape == (20000, 6)
    assert np.allclose(out[:, :3].numpy(), pc[:, :3].numpy(), atol=5e-6)
    radial = (out[:, 3:] * d).sum(-1)
    assert (radial > 0.99).all()
