"""GPU parity of the patch-graph pipeline: util.divide_pc_to_graph
(util.py:160-212) and graph_dipole.graph_dipole_server_api (graph_dipole.py:87-117), against the reference's own
run (tests/golden/patchgraph.npz).  Index work and orientation decisions identical; flip scores within the stated
FP32 tolerance (1e-4 relative to the score + 1e-5 of the largest score); score signs identical."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _csr(parts):
    parts = [p.cpu().numpy() for p in parts]
    return np.concatenate(parts), np.concatenate([[0], np.cumsum([len(p) for p in parts])])


@pytest.mark.parametrize("tag,batched", [("hand", True), ("hand", False), ("fandisk", False)])
def test_divide_pc_to_graph_matches_reference(golden, dev, tag, batched):
    """batched = every patch oriented in one launch (a cluster per patch); otherwise the per-patch calls the
    reference makes.  Both must reproduce the reference's decisions."""
    import dnp_util
    import field_utils
    import graph_dipole
    g = golden("patchgraph")
    pc = torch.from_numpy(g[f"{tag}_pc"]).to(dev)
    estimator = graph_dipole.st_propagation_points_file if tag == "hand" else None
    if estimator is not None and not batched:
        estimator = lambda p, f=estimator: f(p, verbose=False)      # noqa: E731  (a plain callable: per-patch path)
    G, index = dnp_util.divide_pc_to_graph(pc, n_part=int(g[f"{tag}_n_part"]), min_patch=int(g[f"{tag}_min_patch"]),
                                           edge_calculator=field_utils.field_edge_calculator,
                                           point_estimator=estimator)
    perm, off = _csr(index)
    assert np.array_equal(perm, g[f"{tag}_graph_perm"]) and np.array_equal(off, g[f"{tag}_graph_offsets"])
    # per-patch greedy orientation: same decisions, hence the same normals (a 1-point patch turns NaN in both)
    got = pc.cpu().numpy()
    want = g[f"{tag}_graph_pc"]
    assert np.array_equal(np.isnan(got), np.isnan(want))
    assert np.array_equal(np.nan_to_num(got[:, 3:]), np.nan_to_num(want[:, 3:]))
    assert np.allclose(np.nan_to_num(got[:, :3]), np.nan_to_num(want[:, :3]), rtol=0, atol=1e-6)
    if batched:
        return
    # edges and scores
    assert G.V == list(range(len(index)))
    assert np.array_equal(np.array([[e.u, e.v] for e in G.E]), g[f"{tag}_graph_edges"])
    w = np.array([float(e.w) for e in G.E])
    invw = np.array([float(e.invw) for e in G.E])
    w_ref = g[f"{tag}_graph_w"].astype(np.float64)
    assert np.all(np.abs(w - w_ref) <= 1e-4 * np.abs(w_ref) + 1e-5 * np.abs(w_ref).max())
    assert np.array_equal(np.sign(w), np.sign(w_ref))
    assert np.array_equal(invw, -w)
    assert isinstance(G.E[0].w, np.ndarray) and G.E[0].w.dtype == np.float32 and G.E[0].w.ndim == 0
    A, B = G.to_matrix()
    assert np.allclose(A, g[f"{tag}_A"], rtol=1e-4, atol=1e-5 * np.abs(w_ref).max())
    # the same scores through the per-edge path (any callable), as the reference evaluates them
    pc2 = g[f"{tag}_graph_pc"].copy()
    bad = np.isnan(pc2).any(axis=1)                               # 1-point patch: keep its cell, null its dipole
    pc2[bad, :3] = g[f"{tag}_pc"][bad, :3]
    pc2[bad, 3:] = 0
    G2, _ = dnp_util.divide_pc_to_graph(torch.from_numpy(pc2).to(dev),
                                        n_part=int(g[f"{tag}_n_part"]), min_patch=int(g[f"{tag}_min_patch"]),
                                        edge_calculator=lambda s, t: field_utils.field_edge_calculator(s, t))
    w2 = np.array([float(e.w) for e in G2.E])
    keep = np.abs(w_ref) > 0
    assert np.all(np.abs(w2 - w)[keep] <= 2e-4 * np.abs(w)[keep] + 2e-5 * np.abs(w).max())


def test_flip_solve_on_the_reference_graph(golden, dev):
    """The recorded score matrices of the reference run: the solve must return a proven optimum."""
    import graph
    g = golden("patchgraph")
    for tag in ("hand", "fandisk"):
        A, B = g[f"{tag}_A"], g[f"{tag}_B"]
        x = graph.MIQP(A, B)
        assert graph.LAST_STATUS["status"] in (1, 2), tag
        assert graph.cal_loss(x, A, B) == pytest.approx(graph.LAST_STATUS["objective"], rel=1e-12)
        # no single flip improves it
        base = graph.LAST_STATUS["objective"]
        for i in range(len(x)):
            y = x.copy()
            y[i] = 1 - y[i]
            assert graph.cal_loss(y, A, B) <= base + 1e-9 * abs(base)


def test_server_api_orients_a_sphere(dev):
    """graph_dipole_server_api end to end on the wire's input type (float64 xyz only): normals estimated,
    patches oriented, flip scores solved -- the result must be one consistent orientation of the sphere."""
    import graph_dipole
    g = torch.Generator().manual_seed(5)
    d = torch.randn(6000, 3, generator=g, dtype=torch.float64)
    d = d / d.norm(dim=-1, keepdim=True)
    xyz = (d * 0.8 + 0.25).numpy()
    config = {"max_nn": 16, "divide_method": "grid_partition", "n_part": 6, "min_patch": 30}
    out = graph_dipole.graph_dipole_server_api(xyz, config)
    assert isinstance(out, np.ndarray) and out.shape == (6000, 6)
    assert np.allclose(out[:, :3], xyz, atol=1e-5)
    radial = (out[:, 3:] * d.numpy()).sum(-1)
    assert np.abs(np.abs(radial) - 1).max() < 0.05               # PCA normals of a sphere are radial
    assert abs(np.sign(radial).mean()) > 0.999                   # and all point the same way
    assert graph_dipole.graph_dipole_server_api(xyz, {"max_nn": 16, "divide_method": "nope"}) is None


def test_batched_patch_orientation_equals_per_patch_calls(dev):
    """dnp_points_propagate_batched_f32 / dnp_patch_lattice_sign_f32 against the single-cloud entry points on the
    same patches: identical normals (same device code), lattice means within FP32 noise, same flips."""
    import dnp_native
    import field_utils
    g = torch.Generator().manual_seed(77)
    sizes = [1, 2, 37, 512, 1500, 3000, 64, 900]
    parts = []
    for k, n in enumerate(sizes):
        d = torch.randn(n, 3, generator=g)
        d = d / d.norm(dim=-1, keepdim=True)
        pc = torch.cat([d * 0.4 + 0.01 * torch.randn(n, 3, generator=g), d], 1).float()
        pc[torch.rand(n, generator=g) < 0.5, 3:] *= -1
        parts.append(pc)
    work = torch.cat(parts).to(dev)
    offsets = torch.tensor(np.concatenate([[0], np.cumsum(sizes)]), dtype=torch.int64, device=dev)
    single = [p.clone().to(dev) for p in parts]
    for p in single:
        field_utils.strongest_field_propagation_points(p, diffuse=True, starting_point=0)
    dnp_native.points_propagate_batched(work, offsets, max(sizes), diffuse=True, eps=1e-6)
    assert torch.equal(work, torch.cat(single))
    means = dnp_native.patch_lattice_sign(work.clone(), offsets, 10, flip=False).cpu().numpy()
    want = np.array([field_utils.measure_mean_potential(p).item() for p in single])
    assert np.allclose(means, want, rtol=1e-4, atol=1e-5 * np.abs(want).max())
    flipped = work.clone()
    dnp_native.patch_lattice_sign(flipped, offsets, 10, flip=True)
    for k in range(len(sizes)):
        lo, hi = int(offsets[k]), int(offsets[k + 1])
        sign = -1.0 if want[k] < 0 else 1.0
        assert torch.equal(flipped[lo:hi, 3:], work[lo:hi, 3:] * sign)
        assert torch.equal(flipped[lo:hi, :3], work[lo:hi, :3])
