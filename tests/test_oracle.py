"""Pins the CPU oracle (oracle/) to outputs of the UNMODIFIED reference.

The reference has no tests or golden vectors of its own, so the fixtures under
tests/golden/ were produced by running the reference itself on its CPU torch path
(tests/golden/make_golden.py, run in the build container).  Every oracle function is checked here
against them; the GPU parity tests then compare the CUDA path with the oracle and the same fixtures.

Tolerances: the float64 oracle must reproduce the reference's float64 outputs
to ~1e-12 (same formula, different summation order); against the reference's float32 outputs it must
sit inside the reference's own float32 noise floor (max 1.1e-5 per-target relative for E).
Decisions (visit order, flips, final normal signs) must be IDENTICAL.
"""
import numpy as np
import pytest

from conftest import rel_err

F64_TOL = 1e-11      # float64 oracle vs reference run on .double() inputs
F32_TOL = 2e-5       # float64 oracle vs reference float32 outputs (reference's own rounding noise)


def _cases(g):
    return range(int(g["n_cases"]))


# ------------------------------------------------------------------ a1 / a2 leaf functions
def test_potential_matches_reference(golden, orc):
    g = golden("leaf")
    for k in _cases(g):
        src, tgt = g[f"c{k}_src"], g[f"c{k}_tgt"]
        phi = orc.potential(src, tgt)
        assert rel_err(phi, g[f"c{k}_phi64"]) < F64_TOL, k
        assert rel_err(phi, g[f"c{k}_phi32"]) < F32_TOL, k
        phi32 = orc.potential(src, tgt, f64=False)
        assert rel_err(phi32, g[f"c{k}_phi64"]) < F32_TOL, k


def test_field_grad_matches_reference(golden, orc):
    g = golden("leaf")
    for k in _cases(g):
        src, tgt = g[f"c{k}_src"], g[f"c{k}_tgt"]
        E = orc.field_grad(src, tgt)
        assert rel_err(E, g[f"c{k}_E64"]) < F64_TOL, k
        assert rel_err(E, g[f"c{k}_E32"]) < F32_TOL, k
        E6 = orc.field_grad(src, tgt[:, :3], eps=1e-6)
        assert rel_err(E6, g[f"c{k}_E32_eps1e-6"]) < F32_TOL, k
        E32 = orc.field_grad(src, tgt, f64=False)
        assert rel_err(E32, g[f"c{k}_E64"]) < F32_TOL, k


def test_recursive_chunking_is_only_summation_order(golden, orc):
    """The reference's recursive halving (field_utils.py:23-44,73-94, forced here with max_pts=100)
    adds source halves and concatenates target halves: same totals as one un-chunked evaluation."""
    g = golden("leaf")
    E = orc.field_grad(g["rec_src"], g["rec_tgt"])
    phi = orc.potential(g["rec_src"], g["rec_tgt"])
    assert rel_err(E, g["rec_E64"]) < F64_TOL
    assert rel_err(phi, g["rec_phi64"]) < F64_TOL
    assert rel_err(E, g["rec_E32"]) < F32_TOL
    assert rel_err(phi, g["rec_phi32"]) < F32_TOL


def test_coincident_pairs(golden, orc):
    g = golden("leaf")
    cl = g["self_pc"]
    # field_grad drops coincident pairs exactly (field_utils.py:98-106)
    assert rel_err(orc.field_grad(cl, cl), g["self_E64"]) < F64_TOL
    assert rel_err(orc.field_grad(cl, cl), g["self_E32"]) < F32_TOL
    # ... which equals removing the duplicated source
    E = orc.field_grad(g["dup_src"], cl)
    assert rel_err(E, g["dup_E64"]) < F64_TOL
    E_removed = orc.field_grad(g["dup_src"][:-1], cl)
    assert rel_err(E[5], E_removed[5]) < 1e-12
    # potential has no mask: the reference zeroes every target whose leaf total is NaN (field_utils.py:53-54)
    assert np.all(g["self_phi32"] == 0)
    assert np.all(orc.potential(cl, cl, drop_coincident=False) == 0)
    # product semantics (documented divergence): only the coincident pair is dropped
    phi = orc.potential(cl, cl, drop_coincident=True)
    assert np.isfinite(phi).all() and np.abs(phi).max() > 0


def test_analytic_sphere(golden, orc):
    """Uniform outward sphere of N unit dipoles at radius r: phi(centre) = N / r^2."""
    g = golden("leaf")
    sp, probes = g["sphere_pc"], g["sphere_probes"]
    phi = orc.potential(sp, probes)
    assert rel_err(phi, g["sphere_phi64"]) < F64_TOL
    r = np.linalg.norm(sp[:, :3], axis=1).mean()
    assert abs(phi[0] - sp.shape[0] / r ** 2) / phi[0] < 1e-4
    assert abs(phi[2]) < 1e-2 * phi[0]          # far outside the shell the potential is ~0


def test_single_dipole_closed_form(orc):
    src = np.array([[0.1, -0.2, 0.3, 0.0, 0.0, 2.0]], dtype=np.float32)
    tgt = np.array([[0.4, 0.1, -0.5]], dtype=np.float32)
    R = src[0, :3].astype(np.float64) - tgt[0].astype(np.float64)
    r = np.linalg.norm(R)
    n = src[0, 3:].astype(np.float64)
    assert abs(orc.potential(src, tgt)[0] - n @ R / r ** 3) < 1e-12
    E = -(3 * (n @ R / r) * R / r - n) / (r ** 3 + 1e-5)
    assert np.abs(orc.field_grad(src, tgt)[0] - E).max() < 1e-12


# ------------------------------------------------------------------ a3 lattice test
def test_lattice(golden, orc):
    g = golden("lattice")
    assert np.array_equal(orc.gen_grid(10), g["grid"])          # util.py:26-36, bit for bit
    sp = g["sphere_pc"]
    m = orc.measure_mean_potential(sp)
    assert abs(m - float(g["sphere_mean64"])) < 1e-9 * abs(float(g["sphere_mean64"]))
    assert abs(m - float(g["sphere_mean"])) < 1e-4 * abs(float(g["sphere_mean"]))
    assert m > 0                                                   # outward normals: positive inside


# ------------------------------------------------------------------ a4 edge scores
def test_edge_scores(golden, orc):
    g = golden("edge")
    for k in _cases(g):
        S, T = g[f"e{k}_S"], g[f"e{k}_T"]
        w, invw = orc.field_edge_calculator(S, T)
        assert abs(w - float(g[f"e{k}_w64"])) <= 1e-10 * abs(float(g[f"e{k}_w64"])), k
        assert abs(w - float(g[f"e{k}_w"])) <= 1e-4 * abs(float(g[f"e{k}_w"])), k
        assert invw == -w
        assert (1 if w > 0 else -1) == int(g[f"e{k}_bool"][0])
        assert int(g[f"e{k}_count"][0]) == len(S) * len(T)
    cl = g["selfall_pc"]
    w, _ = orc.field_edge_calculator(cl, cl)
    assert abs(w - float(g["selfall_w"])) <= 1e-4 * abs(float(g["selfall_w"]))


# ------------------------------------------------------------------ a5 orientation transfer
def test_reference_field(golden, orc):
    g = golden("transfer")
    out3 = orc.reference_field(g["ref"], g["tgt3"])
    assert np.array_equal(out3[:, :3], g["out3"][:, :3])
    assert np.abs(out3[:, 3:] - g["out3"][:, 3:]).max() < 1e-5
    out6 = orc.reference_field(g["ref"], g["tgt6"])
    assert np.array_equal(out6, g["out6"])                         # signs identical


# ------------------------------------------------------------------ a6 per-point propagation (cfg1, data/ok.xyz)
@pytest.mark.parametrize("tag", ["n512", "n2000", "full"])
def test_points_propagation_bit_exact(golden, orc, tag):
    g = golden("points_prop")
    n = int(g[f"{tag}_n"])
    out, order, flipped = orc.points_propagate(g["pc_in"][:n], diffuse=bool(g[f"{tag}_diffuse"]))
    assert np.array_equal(order, g[f"{tag}_order"])
    assert np.array_equal(flipped, g[f"{tag}_flipped"])
    assert np.array_equal(out[:, 3:], g[f"{tag}_normals"])
    m = orc.measure_mean_potential(out)
    assert abs(m - float(g[f"{tag}_mean_potential"])) <= 1e-4 * abs(float(g[f"{tag}_mean_potential"]))


# ------------------------------------------------------------------ a7 patch propagation (cfg2, data/fandisk.xyz)
def _patches(perm, off):
    return [perm[off[i]:off[i + 1]] for i in range(len(off) - 1)]


@pytest.mark.parametrize("tag,diffuse,use_w", [("plain", False, False), ("diffuse_w", True, True)])
def test_patch_propagation(golden, orc, tag, diffuse, use_w):
    g = golden("patch_prop")
    patches = _patches(g["perm"], g["offsets"])
    kept = [(int(i), patches[int(i)]) for i in g["kept_ids"]]
    out, order = orc.patch_propagate(g["pc_in"], kept, patches, diffuse=diffuse,
                                     weights=g["weights"] if use_w else None)
    assert np.array_equal(order, g[f"{tag}_order"])
    want = g[f"{tag}_normals"]
    assert np.array_equal(np.sign(out[:, 3:]), np.sign(want))      # every final normal sign
    assert np.abs(out[:, 3:] - want).max() < 1e-6


# ------------------------------------------------------------------ a8 representative-point propagation (cfg3 stand-in)
@pytest.mark.parametrize("tag,diffuse", [("plain", False), ("diffuse", True)])
def test_reps_propagation(golden, orc, tag, diffuse):
    g = golden("reps_prop")
    reps = list(zip(_patches(g["rep_perm"], g["rep_off"]), _patches(g["rest_perm"], g["rest_off"])))
    out, order = orc.reps_propagate(g["pc_in"], reps, diffuse=diffuse)
    assert np.array_equal(order, g[f"{tag}_order"])
    assert np.array_equal(out[:, 3:], g[f"{tag}_normals"])


# ------------------------------------------------------------------ the torch port timed as the CPU baseline
def test_torch_port_is_the_reference_op_chain(golden):
    """oracle/torch_port.py keeps the reference's op sequence and chunking, so its float32 outputs equal the
    reference's bit for bit (same ATen kernels, same order)."""
    import torch
    from oracle import torch_port as tp
    g = golden("leaf")
    for k in _cases(g):
        src, tgt = torch.from_numpy(g[f"c{k}_src"]), torch.from_numpy(g[f"c{k}_tgt"])
        assert np.array_equal(tp.potential(src, tgt).numpy(), g[f"c{k}_phi32"]), k
        assert np.array_equal(tp.field_grad(src, tgt).numpy(), g[f"c{k}_E32"]), k
    src, tgt = torch.from_numpy(g["rec_src"]), torch.from_numpy(g["rec_tgt"])
    assert np.array_equal(tp.field_grad(src, tgt, max_pts=100).numpy(), g["rec_E32"])
    assert np.array_equal(tp.potential(src, tgt, max_pts=100).numpy(), g["rec_phi32"])
    cl = torch.from_numpy(g["self_pc"])
    assert np.array_equal(tp.field_grad(cl, cl).numpy(), g["self_E32"])
    assert np.array_equal(tp.potential(cl, cl).numpy(), g["self_phi32"])


# ------------------------------------------------------------------ the "xie" interaction
def test_xie_interaction_matches_reference(golden, orc):
    g = golden("xie")
    src, tgt = g["src"], g["tgt"]
    for C in (3, 2):
        m = orc.xie_intersaction(src, tgt, C=C)
        scale = np.abs(m).max()
        assert np.abs(m - g[f"inter64_C{C}"]).max() <= 1e-12 * scale            # float64 restatement
        assert np.abs(m - g[f"inter_C{C}"]).max() <= 2e-6 * scale              # the reference's own float32 output
    f = orc.xie_field(src, tgt, C=3)
    assert np.abs(f - g["field_C3"]).max() <= 2e-6 * np.abs(f).max()
    # coincident pairs keep the un-normalised source normal (field_utils.py:456-460)
    assert np.allclose(f[np.arange(7), np.arange(7)], src[:7, 3:], rtol=0, atol=1e-7)
    mk = orc.xie_intersaction(src, tgt, C=3, knn_mask=16)
    assert np.array_equal(mk != 0, g["inter_knn16"] != 0)                        # the same kNN mask
    assert np.abs(mk - g["inter_knn16"]).max() <= 2e-6 * np.abs(mk).max()


@pytest.mark.parametrize("diffuse", [False, True])
def test_xie_propagation_in_order_matches_reference(golden, orc, diffuse):
    g = golden("xie")
    flips, _ = orc.xie_propagate_in_order(g["prop_pc"], g["prop_orders"], diffuse=diffuse)
    assert np.array_equal(flips, g[f"prop_flips_f64_diffuse{int(diffuse)}"])
    assert np.array_equal(flips, g[f"prop_flips_f32_diffuse{int(diffuse)}"])


# ------------------------------------------------------------------ voxel partition + merge
def _csr(parts):
    return (np.concatenate(parts) if parts else np.zeros(0, dtype=np.int64),
            np.concatenate([[0], np.cumsum([len(p) for p in parts])]))


@pytest.mark.parametrize("tag", ["fandisk", "ok", "hand"])
def test_partition_matches_reference(golden, orc, tag):
    g = golden("partition")
    idx, ijk = orc.divide_pc_cells(g[f"{tag}_pc"], int(g[f"{tag}_n_part"]))
    perm, off = _csr(idx)
    assert np.array_equal(perm, g[f"{tag}_cell_perm"]) and np.array_equal(off, g[f"{tag}_cell_offsets"])
    assert np.array_equal(ijk, g[f"{tag}_cell_ijk"])
    merged, mijk = orc.merge_nodes(idx, ijk, int(g[f"{tag}_min_patch"]))
    perm, off = _csr(merged)
    assert np.array_equal(perm, g[f"{tag}_merged_perm"]) and np.array_equal(off, g[f"{tag}_merged_offsets"])
    assert np.array_equal(np.concatenate(mijk), g[f"{tag}_merged_ijk"])


@pytest.mark.parametrize("tag", ["fandisk", "ok", "hand"])
def test_fast_merge_equals_the_reference_scan(golden, tag):
    """dnp_util.merge_nodes (cell -> owner map) against the reference's own merge output; host logic, no GPU."""
    import dnp_util
    import torch
    g = golden("partition")
    perm, off, ijk = g[f"{tag}_cell_perm"], g[f"{tag}_cell_offsets"], g[f"{tag}_cell_ijk"]
    idx = [[torch.from_numpy(perm[off[c]:off[c + 1]])] for c in range(len(off) - 1)]
    cells = [[torch.from_numpy(ijk[c])] for c in range(len(ijk))]
    merged, mijk = dnp_util.merge_nodes(None, idx, cells, int(g[f"{tag}_min_patch"]))
    p2, o2 = _csr([m.numpy() for m in merged])
    assert np.array_equal(p2, g[f"{tag}_merged_perm"]) and np.array_equal(o2, g[f"{tag}_merged_offsets"])
    assert np.array_equal(torch.stack([c for sub in mijk for c in sub]).numpy(), g[f"{tag}_merged_ijk"])
    assert np.array_equal(np.concatenate([[0], np.cumsum([len(s) for s in mijk])]), g[f"{tag}_merged_ijk_offsets"])


@pytest.mark.parametrize("tag", ["fandisk", "ok", "hand"])
def test_integer_cell_merge_equals_the_reference_scan(golden, tag):
    """dnp_util.merge_cells (the host half of the fused partition_patches path) against the reference's own merge
    output: same patches, same cell order inside every patch; no GPU."""
    import dnp_util
    g = golden("partition")
    off, ijk = g[f"{tag}_cell_offsets"], g[f"{tag}_cell_ijk"]
    kept, members, size = dnp_util.merge_cells((off[1:] - off[:-1]).tolist(), ijk, int(g[f"{tag}_n_part"]),
                                               int(g[f"{tag}_min_patch"]))
    perm = g[f"{tag}_cell_perm"]
    rows = np.concatenate([perm[off[c]:off[c + 1]] for i in kept for c in members[i]])
    assert np.array_equal(rows, g[f"{tag}_merged_perm"])
    assert np.array_equal(np.concatenate([[0], np.cumsum([size[i] for i in kept])]), g[f"{tag}_merged_offsets"])
    assert np.array_equal(np.stack([ijk[c] for i in kept for c in members[i]]), g[f"{tag}_merged_ijk"])
