"""GPU parity of the voxel partition (util.py:110-150, :448-492): index work, so
bit-exact against the reference's own outputs (tests/golden/partition.npz) and the oracle."""
import inspect
import json
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def csr(parts):
    parts = [np.asarray(p.cpu() if isinstance(p, torch.Tensor) else p) for p in parts]
    return (np.concatenate(parts) if parts else np.zeros(0, dtype=np.int64),
            np.concatenate([[0], np.cumsum([len(p) for p in parts])]))


def test_signatures_match_the_reference():
    import dnp_util
    want = json.load(open(os.path.join(GOLDEN, "partition_signatures.json")))
    for name, params in want.items():
        got = [[p.name, None if p.default is inspect.Parameter.empty else repr(p.default)]
               for p in inspect.signature(getattr(dnp_util, name)).parameters.values()]
        assert got == params, name


@pytest.mark.parametrize("tag", ["fandisk", "ok", "hand"])
@pytest.mark.parametrize("on_gpu", [True, False])
def test_divide_and_merge_golden(golden, dev, tag, on_gpu):
    import dnp_util
    g = golden("partition")
    pc = torch.from_numpy(g[f"{tag}_pc"])
    pc = pc.to(dev) if on_gpu else pc
    n_part, min_patch = int(g[f"{tag}_n_part"]), int(g[f"{tag}_min_patch"])
    idx, ijk = dnp_util._divide_pc(pc[:, :3], n_part, min_patch=min_patch)          # a strided view, as the callers pass
    assert all(len(c) == 1 and c[0].dtype == torch.int64 and c[0].device == pc.device for c in idx)
    perm, off = csr([c[0] for c in idx])
    assert np.array_equal(perm, g[f"{tag}_cell_perm"]) and np.array_equal(off, g[f"{tag}_cell_offsets"])
    assert np.array_equal(torch.stack([c[0] for c in ijk]).cpu().numpy(), g[f"{tag}_cell_ijk"])
    merged, mijk = dnp_util.merge_nodes(pc, idx, ijk, min_patch)
    perm, off = csr(merged)
    assert np.array_equal(perm, g[f"{tag}_merged_perm"]) and np.array_equal(off, g[f"{tag}_merged_offsets"])
    # the fused path the pipelines call (util.divide_pc): the same merged patches and cells, the reference's order
    fast, fast_ijk = dnp_util.partition_patches(pc, n_part, min_patch=min_patch, want_ijk=True)
    perm, off = csr(fast)
    assert np.array_equal(perm, g[f"{tag}_merged_perm"]) and np.array_equal(off, g[f"{tag}_merged_offsets"])
    assert np.array_equal(torch.stack([c for sub in fast_ijk for c in sub]).numpy(), g[f"{tag}_merged_ijk"])
    assert all(p.device == pc.device and p.dtype == torch.int64 for p in fast)
    again = dnp_util.divide_pc(pc, n_part, min_patch=min_patch)
    assert len(again) == len(fast) and all(torch.equal(a, b) for a, b in zip(again, fast))


@pytest.mark.parametrize("n,n_part", [(1, 1), (7, 2), (1000, 3), (50000, 41), (300000, 255)])
def test_sizes_vs_oracle(orc, dev, n, n_part):
    import dnp_native as dn
    g = torch.Generator().manual_seed(n + n_part)
    pc = (torch.rand(n, 3, generator=g) - 0.5) * 3.2                                # some points outside (-1.5, 1.5]
    edge = 3.0 / n_part
    pc[::5] = torch.round(pc[::5] / edge) * edge                                     # many points exactly on cell faces
    if n > 10:
        pc[3, 0] = float("nan")
        pc[4, 1] = float("inf")
        pc[6] = torch.tensor([-1.5, 0.0, 0.0])                                       # the open lower face of the box
    perm, offsets, ijk = dn.voxel_partition(pc.to(dev), n_part)
    want_idx, want_ijk = orc.divide_pc_cells(pc.numpy(), n_part)
    wp, wo = csr(want_idx)
    assert np.array_equal(perm.cpu().numpy(), wp)
    assert np.array_equal(offsets.cpu().numpy(), wo)
    assert np.array_equal(ijk.cpu().numpy(), want_ijk)


def test_full_size_properties(dev):
    """1 048 576 points at --number_parts 41 (the lion demo's flag): every in-box point lands in exactly one cell,
    cells are sorted, rows ascend inside a cell, and every row satisfies its cell's bounds."""
    import dnp_native as dn
    n, n_part = 1 << 20, 41
    g = torch.Generator().manual_seed(1)
    d = torch.randn(n, 3, generator=g)
    pc = (d / d.norm(dim=-1, keepdim=True) * 0.45 + torch.randn(n, 3, generator=g) * 0.002).float().to(dev)
    perm, offsets, ijk = dn.voxel_partition(pc, n_part)
    assert perm.numel() == n and int(offsets[-1]) == n
    assert torch.equal(torch.sort(perm).values, torch.arange(n, device=dev))
    key = (ijk[:, 0].long() * 42 + ijk[:, 1]) * 42 + ijk[:, 2]
    assert bool((key[1:] > key[:-1]).all())
    sizes = offsets[1:] - offsets[:-1]
    assert bool((sizes > 0).all())
    cell_of_row = torch.repeat_interleave(torch.arange(ijk.shape[0], device=dev), sizes)
    same = cell_of_row[1:] == cell_of_row[:-1]
    assert bool((perm[1:][same] > perm[:-1][same]).all())
    edge = 3.0 / n_part
    lo = (ijk[cell_of_row].double() * edge - 1.5).float()
    hi = (ijk[cell_of_row].double() * edge - 1.5 + edge).float()
    x = pc[perm]
    assert bool(((lo < x) & (x <= hi)).all())


def test_install_points_the_reference_util_at_the_gpu_partition():
    import types
    import dnp_util
    fake = types.ModuleType("util")
    dnp_util.install(fake)
    assert fake._divide_pc is dnp_util._divide_pc and fake.merge_nodes is dnp_util.merge_nodes


def test_errors(dev):
    import dnp_native as dn
    import dnp_util
    with pytest.raises(dn.DipoleError):
        dn.voxel_partition(torch.zeros(10, 3, device=dev), 256)
    with pytest.raises(TypeError):
        dnp_util._divide_pc(torch.zeros(10, 3, dtype=torch.float64, device=dev), 4)
    idx, ijk = dnp_util._divide_pc(torch.zeros(0, 3, device=dev), 4)
    assert idx == [] and ijk == []
