"""The algebra behind the slab schedule of the greedy patch loop (dipole-normal-prop_b200/csrc/dnp_slab.cu), on the CPU.

The schedule never runs the reference's loop (field_utils.py:239-265): it computes the field of every patch with the
normals AS GIVEN (one slab per patch), reduces the slabs to a P x P matrix W[s][p] = sum_{t in p} slab_s(t) . n_t,
takes the greedy decisions on W alone (I += sigma_s W[s][:], first abs-arg-max, sigma = sign), and assembles E from the
slabs in visit order.  That is legitimate because the field is linear and exactly odd in the source normals.  This file
restates those four phases in numpy on top of the oracle's field_grad and pins them to the decisions the UNMODIFIED
reference recorded (tests/golden/reps_prop.npz, patch_prop.npz) -- so the identity is checked here without a GPU, and
the GPU tests only have to show that the kernels implement this restatement."""
import numpy as np
import pytest


def slab_schedule(orc, pts, patches, start, diffuse, f32, symmetric=False):
    """Phases A-D of dnp_slab.cu.  patches: index arrays (disjoint, the targets are their union in CSR order).
    Returns (visit order, sigma per patch, E over the CSR rows)."""
    P = len(patches)
    perm = np.concatenate(patches)
    offsets = np.concatenate([[0], np.cumsum([len(p) for p in patches])])
    patch_of = np.repeat(np.arange(P), [len(p) for p in patches])
    tgt = pts[perm]
    dt = np.float32 if f32 else np.float64
    # A: one slab per patch, normals as on entry
    slabs = np.stack([orc.field_grad(pts[p], tgt, f64=not f32).astype(dt) if len(p) else np.zeros((len(perm), 3), dt)
                      for p in patches])
    # B: the interaction matrix
    nrm = tgt[:, 3:].astype(dt)
    dots = (slabs * nrm[None]).sum(-1)                                            # [P, T]
    W = np.stack([dots[:, offsets[p]:offsets[p + 1]].sum(-1) for p in range(P)], axis=1).astype(dt)
    if symmetric:       # every patch pair once (s < p), the other half copied: the interaction energy is symmetric
        W = np.triu(W, 1) + np.triu(W, 1).T
    # C: the greedy loop on W (field_utils.py:244-252 / :316-323)
    inter = np.zeros(P, dt)
    active = np.ones(P, bool)
    sigma = np.ones(P, dt)
    rank = np.full(P, P)
    order = []
    for step in range(P):
        if step == 0:
            idx, val = start, 0.0
        else:
            cand = np.where(active, np.abs(inter), -1.0)
            idx = int(np.argmax(cand))                                            # first maximum, like torch
            val = inter[idx]
        order.append(idx)
        active[idx] = False
        sigma[idx] = -1.0 if val < 0 else 1.0
        rank[idx] = step
        inter[active] += sigma[idx] * W[idx, active]
    # D: E in visit order (own patch skipped; not diffuse: only the visits before the target's own patch)
    E = np.zeros((len(perm), 3), dt)
    for k, s in enumerate(order):
        take = patch_of != s
        if not diffuse:
            take &= rank[patch_of] > k
        E[take] += sigma[s] * slabs[s][take]
    return order, sigma, E, perm, patch_of


def _patches(perm, off):
    return [perm[off[i]:off[i + 1]].copy() for i in range(len(off) - 1)]


@pytest.mark.parametrize("f32", [False, True])
@pytest.mark.parametrize("tag,diffuse", [("plain", False), ("diffuse", True)])
def test_matrix_form_takes_the_reference_decisions_reps(golden, orc, tag, diffuse, f32):
    g = golden("reps_prop")
    pts = g["pc_in"].astype(np.float32)
    reps = _patches(g["rep_perm"], g["rep_off"])
    rests = _patches(g["rest_perm"], g["rest_off"])
    want_order = list(g[f"{tag}_order"])
    order, sigma, E, perm, patch_of = slab_schedule(orc, pts, reps, want_order[0], diffuse, f32)
    assert order == want_order                                                    # every greedy decision
    # representatives: patch flip (sigma), then -- when diffuse -- the per-point sign of E.n (:267-271)
    got = pts.copy()
    for p, (rep, rest) in enumerate(zip(reps, rests)):
        got[rep, 3:] *= sigma[p]
        got[rest, 3:] *= sigma[p]
    if diffuse:
        dot = (E * got[perm, 3:]).sum(-1)
        got[perm, 3:] *= np.where(dot > 0, 1.0, -1.0)[:, None]
    rep_rows = np.concatenate(reps)
    assert np.array_equal(np.sign(got[rep_rows, 3:]), np.sign(g[f"{tag}_normals"][rep_rows]))


@pytest.mark.parametrize("tag,diffuse", [("plain", False)])
def test_matrix_form_takes_the_reference_decisions_all_points(golden, orc, tag, diffuse):
    """fandisk, 72 patches covering the cloud (field_utils.py:286-348)."""
    g = golden("patch_prop")
    pts = g["pc_in"].astype(np.float32)
    patches = _patches(g["perm"], g["offsets"])
    want_order = list(g[f"{tag}_order"])
    order, sigma, E, perm, patch_of = slab_schedule(orc, pts, patches, want_order[0], diffuse, f32=False)
    assert order == want_order
    got = pts.copy()
    for p, rows in enumerate(patches):
        got[rows, 3:] *= sigma[p]
    assert np.array_equal(np.sign(got[:, 3:]), np.sign(g[f"{tag}_normals"]))


def test_running_field_equals_the_loops(golden, orc):
    """E assembled from the slabs = the running field of the oracle's restatement of the reference loop."""
    g = golden("reps_prop")
    pts = g["pc_in"].astype(np.float32)
    reps = _patches(g["rep_perm"], g["rep_off"])
    for diffuse, tag in ((True, "diffuse"), (False, "plain")):
        order, sigma, E, perm, patch_of = slab_schedule(orc, pts, reps, int(g[f"{tag}_order"][0]), diffuse, f32=False)
        # the loop, stated directly
        work = pts.copy()
        E_loop = np.zeros((len(pts), 3))
        visited = np.zeros(len(reps), bool)
        for k, s in enumerate(order):
            work[reps[s], 3:] *= sigma[s]
            visited[s] = True
            mask = np.zeros(len(pts), bool)
            if diffuse:
                mask[perm] = True
                mask[reps[s]] = False
            else:
                for p in np.nonzero(~visited)[0]:
                    mask[reps[p]] = True
            if mask.any():
                E_loop[mask] += orc.field_grad(work[reps[s]], work[mask])
        assert np.abs(E - E_loop[perm]).max() <= 1e-10 * np.abs(E_loop).max()


@pytest.mark.parametrize("f32", [False, True])
def test_symmetric_form_takes_the_reference_decisions(golden, orc, f32):
    """Not diffuse, the schedule computes every patch PAIR once (patch s acts on the patches after it) and mirrors the
    matrix: n_j . E_i(q_j) = w [n_i . n_j - 3 (n_i . R)(n_j . R) / r^2] is symmetric in i and j.  Same decisions as the
    reference on both fixtures, and the two halves of the full matrix agree to rounding."""
    g = golden("reps_prop")
    pts = g["pc_in"].astype(np.float32)
    reps = _patches(g["rep_perm"], g["rep_off"])
    want_order = list(g["plain_order"])
    order, sigma, _, _, _ = slab_schedule(orc, pts, reps, want_order[0], False, f32, symmetric=True)
    assert order == want_order
    g2 = golden("patch_prop")
    pts2 = g2["pc_in"].astype(np.float32)
    patches = _patches(g2["perm"], g2["offsets"])
    want2 = list(g2["plain_order"])
    order2, sigma2, _, _, _ = slab_schedule(orc, pts2, patches, want2[0], False, f32, symmetric=True)
    assert order2 == want2
    got = pts2.copy()
    for p, rows in enumerate(patches):
        got[rows, 3:] *= sigma2[p]
    assert np.array_equal(np.sign(got[:, 3:]), np.sign(g2["plain_normals"]))


def test_interaction_matrix_is_symmetric(golden, orc):
    g = golden("reps_prop")
    pts = g["pc_in"].astype(np.float32)
    reps = _patches(g["rep_perm"], g["rep_off"])
    perm = np.concatenate(reps)
    off = np.concatenate([[0], np.cumsum([len(p) for p in reps])])
    tgt = pts[perm]
    W = np.zeros((len(reps), len(reps)))
    for s, rows in enumerate(reps):
        dots = (orc.field_grad(pts[rows], tgt) * tgt[:, 3:]).sum(-1)
        W[s] = [dots[off[p]:off[p + 1]].sum() for p in range(len(reps))]
    np.fill_diagonal(W, 0.0)
    assert np.abs(W - W.T).max() <= 1e-12 * np.abs(W).max()
