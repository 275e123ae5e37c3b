"""Pins oracle/torch_drivers.py (the reference's four drivers restated with the reference's ATen op sequence; the
CPU side of bench.py's "e2e orient sec") to the fixtures the UNMODIFIED reference generated
(tests/golden/make_golden.py).  Same ops in the same order => the float32 normals are the reference's bit for bit,
and so are the greedy decisions.  No GPU."""
import numpy as np
import pytest
import torch

from oracle import torch_drivers as td


def _patches(perm, off):
    return [torch.from_numpy(perm[off[i]:off[i + 1]].copy()) for i in range(len(off) - 1)]


@pytest.mark.parametrize("tag", ["n512", "n2000"])
def test_point_propagation_is_the_reference_bit_for_bit(golden, tag):
    """field_utils.py:353-388 on data/ok.xyz (cfg1): visit order, flips, final normals, lattice mean."""
    g = golden("points_prop")
    n, diffuse = int(g[f"{tag}_n"]), bool(g[f"{tag}_diffuse"])
    pts = torch.from_numpy(g["pc_in"][:n].copy())
    trace, flips = [], []
    out = td.propagate_points(pts, diffuse=diffuse, starting_point=0, trace=trace, flips=flips)
    assert np.array_equal(np.array(trace, dtype=np.int32), g[f"{tag}_order"])
    assert np.array_equal(np.array([False] + flips), g[f"{tag}_flipped"])
    assert np.array_equal(out[:, 3:].numpy(), g[f"{tag}_normals"])
    assert np.array_equal(td.lattice_mean_potential(out).numpy(), g[f"{tag}_mean_potential"])


def test_point_propagation_prefix_of_the_full_cloud(golden):
    """The bounded sample bench.py times for cfg1 (first steps of the full N = 10 000 run) follows the golden order."""
    g = golden("points_prop")
    pts = torch.from_numpy(g["pc_in"].copy())
    trace = []
    td.propagate_points(pts, trace=trace, max_steps=40)
    assert np.array_equal(np.array(trace, dtype=np.int32), g["full_order"][:41])


@pytest.mark.parametrize("tag,diffuse,use_w", [("plain", False, False), ("diffuse_w", True, True)])
def test_patch_propagation_is_the_reference_bit_for_bit(golden, tag, diffuse, use_w):
    """field_utils.py:286-348 on data/fandisk.xyz, P = 72 (cfg2)."""
    g = golden("patch_prop")
    groups = _patches(g["perm"], g["offsets"])
    kept = [(int(i), groups[int(i)]) for i in g["kept_ids"]]
    pts = torch.from_numpy(g["pc_in"].copy())
    trace = []
    td.propagate_patches(pts, kept, groups, diffuse=diffuse,
                         weights=torch.from_numpy(g["weights"].copy()) if use_w else None, trace=trace)
    assert np.array_equal(np.array(trace, dtype=np.int32), g[f"{tag}_order"])
    assert np.array_equal(pts[:, 3:].numpy(), g[f"{tag}_normals"])
    assert np.array_equal(td.lattice_mean_potential(pts).numpy(), g[f"{tag}_mean_potential"])


@pytest.mark.parametrize("tag,diffuse", [("plain", False), ("diffuse", True)])
def test_reps_propagation_is_the_reference_bit_for_bit(golden, tag, diffuse):
    """field_utils.py:207-282 on the P = 16 stand-in (cfg3)."""
    g = golden("reps_prop")
    reps = list(zip(_patches(g["rep_perm"], g["rep_off"]), _patches(g["rest_perm"], g["rest_off"])))
    pts = torch.from_numpy(g["pc_in"].copy())
    trace, stages = [], {}
    td.propagate_representatives(pts, reps, diffuse=diffuse, trace=trace, stage_seconds=stages)
    assert np.array_equal(np.array(trace, dtype=np.int32), g[f"{tag}_order"])
    assert np.array_equal(pts[:, 3:].numpy(), g[f"{tag}_normals"])
    assert set(stages) == {"propagation", "final_field"}


def test_orientation_transfer_is_the_reference_bit_for_bit(golden):
    """field_utils.py:188-201 (cfg4): 3-column targets get E/|E| appended, 6-column targets are flipped in place."""
    g = golden("transfer")
    ref = torch.from_numpy(g["ref"].copy())
    out3 = td.transfer_orientation(ref, torch.from_numpy(g["tgt3"].copy()))
    assert np.array_equal(out3.numpy(), g["out3"])
    tgt6 = torch.from_numpy(g["tgt6"].copy())
    out6 = td.transfer_orientation(ref, tgt6)
    assert out6 is tgt6 and np.array_equal(out6.numpy(), g["out6"])
