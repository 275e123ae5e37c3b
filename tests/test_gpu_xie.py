"""GPU parity of the "xie" reflected-normal interaction (field_utils.py:431-469,
:509-519, :569-605) through the reference-facing mirror -> ctypes -> libdipole.so.

Tolerance: an entry of the interaction matrix is (n_s - C (n_s.Rh) Rh).n_t / |R|^3, whose numerator cancels, so
errors are measured against the entry's natural scale (1 + C) |n_s| |n_t| / |R|^3:  |got - ref| <= 1e-5 scale.
Flip decisions of the in-order propagation must be identical to the reference's on the golden cloud."""
import json
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def fu():
    import field_utils
    return field_utils


def entry_scale(src, tgt, C):
    R = src[None, :, :3].astype(np.float64) - tgt[:, None, :3]
    r = np.sqrt((R * R).sum(-1))
    ns = np.linalg.norm(src[:, 3:], axis=1)[None, :]
    nt = np.linalg.norm(tgt[:, 3:6], axis=1)[:, None] if tgt.shape[1] >= 6 else 1.0
    return np.where(r == 0, ns * nt, (1 + C) * ns * nt / np.where(r == 0, 1, r) ** 3)


def test_signatures_match_the_reference(fu):
    import inspect
    want = json.load(open(os.path.join(GOLDEN, "xie_signatures.json")))
    for name, params in want.items():
        got = [[p.name, None if p.default is inspect.Parameter.empty else repr(p.default)]
               for p in inspect.signature(getattr(fu, name)).parameters.values()]
        assert got == params, name


def test_interaction_and_field_golden(fu, golden, dev):
    g = golden("xie")
    src, tgt = torch.from_numpy(g["src"]).to(dev), torch.from_numpy(g["tgt"]).to(dev)
    for C in (3, 2):
        m = fu.xie_intersaction(src, tgt, 1e-2, -1, C)
        assert m.shape == (120, 200) and m.is_cuda and m.dtype == torch.float32
        err = np.abs(m.cpu().numpy() - g[f"inter64_C{C}"])
        assert np.all(err <= 1e-5 * entry_scale(g["src"], g["tgt"], C))
    f = fu.xie_field(src, tgt, 1e-2, C=3).cpu().numpy()
    assert f.shape == (120, 200, 3)
    assert np.all(np.abs(f - g["field_C3"]).max(-1) <= 1e-5 * entry_scale(g["src"], g["tgt"][:, :3], 3))
    assert np.array_equal(f[np.arange(7), np.arange(7)], g["src"][:7, 3:])        # coincident pairs keep n_s exactly
    # CPU tensors in -> CPU tensor out, 3-column targets are enough for the field
    f2 = fu.xie_field(torch.from_numpy(g["src"]), torch.from_numpy(g["tgt"][:, :3].copy()), 1e-2, C=3)
    assert not f2.is_cuda and np.array_equal(f2.numpy(), f)


def test_knn_mask_golden(fu, golden, orc, dev):
    g = golden("xie")
    src, tgt = torch.from_numpy(g["src"]).to(dev), torch.from_numpy(g["tgt"]).to(dev)
    m = fu.xie_intersaction(src, tgt, 1e-2, 16, 3).cpu().numpy()
    assert np.array_equal(m != 0, g["inter_knn16"] != 0)                          # same mask as the reference's KD-tree
    assert np.all(np.abs(m - g["inter_knn16"]) <= 1e-5 * entry_scale(g["src"], g["tgt"], 3))
    assert ((m != 0).sum(0) == 16).all()                                          # every source reaches exactly k targets


@pytest.mark.parametrize("S,M", [(1, 1), (5, 3), (129, 65), (1000, 777), (4097, 3000)])
def test_matrix_sizes_vs_oracle(fu, orc, dev, S, M):
    g = torch.Generator().manual_seed(S * 31 + M)
    def cloud(n):
        d = torch.randn(n, 3, generator=g)
        d = d / d.norm(dim=-1, keepdim=True)
        return torch.cat([d * 0.45 + torch.randn(n, 3, generator=g) * 0.01, d * (0.5 + torch.rand(n, 1, generator=g))], 1).float()
    src, tgt = cloud(S), cloud(M)
    m = fu.xie_intersaction(src.to(dev), tgt.to(dev), 1e-2, -1, 2.5).cpu().numpy()
    ref = orc.xie_intersaction(src.numpy(), tgt.numpy(), C=2.5)
    assert np.all(np.abs(m - ref) <= 1e-5 * entry_scale(src.numpy(), tgt.numpy(), 2.5))


@pytest.mark.parametrize("diffuse", [False, True])
def test_propagation_in_order_golden(fu, golden, dev, diffuse):
    g = golden("xie")
    pc = torch.from_numpy(g["prop_pc"]).to(dev)
    flips = fu.xie_propagation_points_in_order(pc, 1e-2, g["prop_orders"], diffuse=diffuse, knn_mask=-1, C=3)
    assert flips.dtype == torch.bool and flips.shape == (5, 1500) and flips.is_cuda
    assert np.array_equal(flips.cpu().numpy(), g[f"prop_flips_f32_diffuse{int(diffuse)}"])
    assert np.array_equal(pc.cpu().numpy(), g["prop_pc"])                        # the reference works on a clone
    # a CPU tensor comes back on the CPU
    cpu = fu.xie_propagation_points_in_order(torch.from_numpy(g["prop_pc"]), 1e-2, g["prop_orders"][:2], diffuse=diffuse)
    assert not cpu.is_cuda and np.array_equal(cpu.numpy(), g[f"prop_flips_f32_diffuse{int(diffuse)}"][:2])


@pytest.mark.parametrize("n,T", [(1, 1), (2, 3), (33, 2), (2049, 9), (6000, 3)])
def test_propagation_sizes_vs_oracle(orc, dev, n, T):
    """Cluster sizes 1..16 CTAs, ragged tails, more chains than one pass of the diffuse kernel handles."""
    import dnp_native as dn
    g = torch.Generator().manual_seed(7 * n + T)
    d = torch.randn(n, 3, generator=g)
    d = d / d.norm(dim=-1, keepdim=True)
    pc = torch.cat([d * 0.45 + torch.randn(n, 3, generator=g) * 0.01, d], 1).float()
    pc[torch.rand(n, generator=g) < 0.5, 3:] *= -1
    orders = torch.stack([torch.randperm(n, generator=g) for _ in range(T)])
    for diffuse in (False, True):
        want_flips, want_inter = orc.xie_propagate_in_order(pc.numpy(), orders.numpy(), diffuse=diffuse)
        flips, inter = dn.xie_propagate_in_order(pc.to(dev), orders, diffuse=diffuse)
        inter = inter.cpu().numpy()
        scale = np.abs(want_inter).max() + 1e-30
        assert np.all(np.abs(inter - want_inter) <= 1e-4 * np.abs(want_inter) + 1e-6 * scale)
        differ = flips.cpu().numpy() != want_flips
        # a decision may only differ where the interaction is within rounding of zero (and then the chain diverges)
        assert not differ.any(), f"{differ.sum()} decisions differ (n={n}, T={T}, diffuse={diffuse})"


def test_propagation_knn_mask_vs_oracle(orc, dev):
    import dnp_native as dn
    g = torch.Generator().manual_seed(5)
    n = 1200
    d = torch.randn(n, 3, generator=g)
    d = d / d.norm(dim=-1, keepdim=True)
    pc = torch.cat([d * 0.45 + torch.randn(n, 3, generator=g) * 0.01, d], 1).float()
    pc[torch.rand(n, generator=g) < 0.5, 3:] *= -1
    orders = torch.stack([torch.randperm(n, generator=g) for _ in range(3)])
    want, _ = orc.xie_propagate_in_order(pc.numpy(), orders.numpy(), diffuse=True, knn_mask=40)
    pd = pc.to(dev)
    got, _ = dn.xie_propagate_in_order(pd, orders, radius2=dn.knn_radius2(pd, pd, 40), diffuse=True)
    assert np.array_equal(got.cpu().numpy(), want)


def test_errors(fu, dev):
    import dnp_native as dn
    pc = torch.zeros(10, 6, device=dev)
    with pytest.raises(ValueError):
        dn.xie_propagate_in_order(pc, torch.zeros(2, 9, dtype=torch.int64))
    with pytest.raises(ValueError):
        dn.xie_propagate_in_order(pc, torch.full((1, 10), 10, dtype=torch.int64))
    with pytest.raises(TypeError):
        fu.xie_intersaction(pc.double(), pc.double(), 1e-2, -1, 3)
    big = int(dn.load().dnp_xie_max_n(0)) + 1
    with pytest.raises(dn.DipoleError):
        dn.xie_propagate_in_order(torch.zeros(big, 6, device=dev), torch.arange(big)[None])


@pytest.mark.parametrize("S,M,k", [(1, 1, 1), (300, 257, 16), (5000, 7001, 30), (2049, 1024, 64)])
def test_knn_radius_kernel(dev, S, M, k):
    """dnp_knn_radius2_f64 (the kNN mask of field_utils.py:447-455 as one radius per source) against a float64
    all-distances + sort: exact equality (both take exact differences of the float coordinates in double)."""
    import dnp_native as dn
    g = torch.Generator().manual_seed(S + M + k)
    src, tgt = torch.rand(S, 6, generator=g).float().to(dev), torch.rand(M, 3, generator=g).float().to(dev)
    if M > 10:
        tgt[5] = tgt[4]                                             # duplicated targets: ties at the k-th distance
    got = dn.knn_radius2(src, tgt, k)
    d = src[:, None, :3].double() - tgt[None, :, :3].double()
    d2 = (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2]
    want = torch.sort(d2, dim=1).values[:, min(k, M) - 1]
    assert torch.equal(got, want)
    wide = torch.zeros(S, 9, device=dev)
    wide[:, 2:8] = src
    assert torch.equal(dn.knn_radius2(wide[:, 2:8], tgt, k), want)   # strided view
