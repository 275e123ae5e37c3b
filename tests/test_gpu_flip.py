"""GPU parity of the patch-graph flip solve (graph.py:431-470, field_utils.py:606-709).

Oracle: exhaustive enumeration of the reference's objective (oracle.miqp_bruteforce; cal_loss itself is pinned to
the reference in tests/test_graph.py).  Bar: the objective equals the optimum to 1e-12 relative (it is evaluated in
double from A and B), and the assignment equals the oracle's up to the global flip x -> 1 - x, which never changes
the objective (Gurobi returns either)."""
import inspect
import json
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def solve(A, B, dev, **kw):
    import dnp_native
    x, obj, status = dnp_native.flip_solve(torch.from_numpy(np.ascontiguousarray(A)).to(dev),
                                           torch.from_numpy(np.ascontiguousarray(B)).to(dev), **kw)
    obj = obj.cpu().numpy()
    return x.cpu().numpy().astype(np.float64), float(obj[0]), float(obj[1]), int(status.item())


def frustrated(rng, n, density=0.5):
    W = np.triu(rng.normal(size=(n, n)), 1) * (rng.random((n, n)) < density)
    return W + W.T


def planted(rng, n, density, flips=0.4):
    """Unfrustrated patch graph: every edge agrees with one hidden orientation."""
    state = (rng.random(n) < flips).astype(np.int64)
    W = np.triu(rng.uniform(0.1, 1.0, (n, n)), 1) * (rng.random((n, n)) < density)
    W[np.arange(n - 1), np.arange(1, n)] = rng.uniform(0.1, 1.0, n - 1)          # connected
    A = (W + W.T) * np.where(state[:, None] == state[None, :], 1.0, -1.0)
    return A, state


def canonical(x):
    x = np.asarray(x, dtype=np.int64)
    n, ones = len(x), int(x.sum())
    return 1 - x if (2 * ones > n or (2 * ones == n and x[0] == 1)) else x


@pytest.mark.parametrize("n", [1, 2, 3, 5, 9, 10, 11, 12, 13, 17, 20])
def test_exhaustive_equals_enumeration(orc, dev, n):
    rng = np.random.default_rng(100 + n)
    A = frustrated(rng, n, 0.7) if n > 1 else np.array([[0.75]])
    want_x, want_obj, n_opt = orc.miqp_bruteforce(A, -A, count_optimal=True)
    x, obj, bound, status = solve(A, -A, dev)
    assert status == 1
    assert obj == pytest.approx(want_obj, rel=1e-12, abs=1e-12)
    assert obj == pytest.approx(orc.cal_loss(x, A, -A), rel=1e-12, abs=1e-12)
    assert bound >= obj - 1e-9
    assert np.array_equal(x, canonical(x))                       # fewer flips of the two twins
    if n_opt <= 2:
        assert np.array_equal(x, canonical(want_x))


@pytest.mark.parametrize("tag", ["sym5", "sym12", "gen9"])
def test_reference_matrices(golden, orc, dev, tag):
    """The recorded matrices, incl. the general case (A, B unrelated, not symmetric, non-zero diagonal)."""
    g = golden("graph")
    A, B = g[f"{tag}_A"], g[f"{tag}_B"]
    want_x, want_obj = orc.miqp_bruteforce(A, B)
    x, obj, bound, status = solve(A, B, dev)
    assert obj == pytest.approx(want_obj, rel=1e-12)
    assert obj >= g[f"{tag}_loss"].max()
    assert np.array_equal(x, canonical(want_x))


def test_exhaustive_large_and_anneal_agree(dev):
    """n = 30 (2^29 states walked): the planted instance has a known unique optimum; on a frustrated instance
    the annealing path (forced with exact_max_n = 0) must reach the exhaustive optimum."""
    rng = np.random.default_rng(7)
    A, state = planted(rng, 30, 0.3)
    x, obj, bound, status = solve(A, -A, dev)
    assert status == 1 and np.array_equal(x, canonical(state))
    assert obj == pytest.approx(np.abs(A).sum(), rel=1e-12) and obj == pytest.approx(bound, rel=1e-12)
    A = frustrated(rng, 28, 0.4)
    x1, obj1, _, st1 = solve(A, -A, dev)
    x2, obj2, _, st2 = solve(A, -A, dev, exact_max_n=0)
    assert st1 == 1 and st2 in (0, 2)
    assert obj2 == pytest.approx(obj1, rel=1e-12)
    assert np.array_equal(x1, x2)


def test_anneal_certifies_unfrustrated_graphs(dev):
    """Patch graphs of a consistently orientable surface are unfrustrated: the bound certificate makes the
    heuristic path exact there.  Sparse (grid-like) and dense variants, up to the 4096-patch limit."""
    rng = np.random.default_rng(11)
    for n, density in ((72, 0.08), (300, 0.02), (1000, 0.004), (4096, 0.0005)):
        A, state = planted(rng, n, density)
        x, obj, bound, status = solve(A, -A, dev, sweeps=256)
        assert status == 2, n
        assert np.array_equal(x, canonical(state)), n
        assert obj == pytest.approx(np.abs(A).sum(), rel=1e-12)


def test_anneal_frustrated_is_local_optimum_and_deterministic(dev):
    rng = np.random.default_rng(13)
    n = 200
    A = frustrated(rng, n, 0.05)
    x, obj, bound, status = solve(A, -A, dev, seed=3)
    x_again, obj_again, _, _ = solve(A, -A, dev, seed=3)
    assert np.array_equal(x, x_again) and obj == obj_again
    assert status == 0 and obj < bound
    s = 1.0 - 2.0 * x
    gain = -4.0 * s * (A @ s)                                     # objective change of each single flip
    assert gain.max() <= 1e-9 * np.abs(A).sum()
    # better than plain descent from "no flip"
    s0 = np.ones(n)
    for _ in range(50):
        for i in range(n):
            if -s0[i] * (A[i] @ s0) > 0:
                s0[i] = -s0[i]
    assert obj >= float(s0 @ A @ s0) - 1e-9


def test_miqp_api(golden, orc, dev, capsys):
    import field_utils
    import graph
    g = golden("graph")
    A, B = g["sym12_A"], g["sym12_B"]
    x = graph.MIQP(A, B)
    assert isinstance(x, np.ndarray) and x.dtype == np.float64 and x.shape == (12,)
    want_x, want_obj = orc.miqp_bruteforce(A, B)
    assert "Obj: %g" % want_obj in capsys.readouterr().out
    assert graph.cal_loss(x, A, B) == pytest.approx(want_obj, rel=1e-12)
    assert graph.LAST_STATUS["status"] == 1
    # field_utils.MIQP minimises (objective -cal_loss, field_utils.py:628-629)
    y = field_utils.MIQP(A, B)
    lo_x, lo_obj = orc.miqp_bruteforce(-A, -B)
    assert field_utils.cal_loss(y, A, B) == pytest.approx(-lo_obj, rel=1e-12)
    want = json.load(open(os.path.join(GOLDEN, "graph_field_utils_signatures.json")))
    for name, params in want.items():
        got = [[p.name, None if p.default is inspect.Parameter.empty else repr(p.default)]
               for p in inspect.signature(getattr(field_utils, name)).parameters.values()]
        assert got == params, name


@pytest.mark.parametrize("times", [1, 5])
def test_bfstree_vote_matches_reference(golden, dev, times):
    """field_utils.xie_propagation_points_onbfstree against the reference's own run (Gurobi replaced by
    enumeration in the generator): same vote matrices imply the same chains; final normals equal up to the
    global sign the tie between x and 1 - x leaves open."""
    import field_utils
    g = golden("graph")
    pts = torch.from_numpy(g["tree_pc"]).to(dev)
    work = pts.clone()
    mask = field_utils.xie_propagation_points_onbfstree(work, 1e-2, diffuse=True, starting_point=0, times=times, k=10,
                                                        treshold=0.1)
    assert mask.dtype == torch.bool and mask.shape == (pts.shape[0],) and mask.device == work.device     # :709
    assert torch.equal(work[mask, 3:], -pts[mask, 3:]) and torch.equal(work[~mask, 3:], pts[~mask, 3:])
    got = work.cpu().numpy()
    want = g[f"tree_result_t{times}"]
    assert np.array_equal(got[:, :3], want[:, :3])
    same = np.array_equal(got[:, 3:], want[:, 3:])
    twin = np.array_equal(got[:, 3:], -want[:, 3:])              # the other optimum of the vote: every chain inverted
    assert same or twin
