"""CPU tests of the flip-solve row: the oracle's cal_loss against values recorded
from the reference's graph.cal_loss (tests/golden/graph.npz), the enumeration oracle's own consistency, and the
host-side mirror of graph.py (container, BFS routes, signatures).  The maximiser itself is Gurobi's in the
reference (absent): the oracle for it is exhaustive enumeration of the pinned objective."""
import inspect
import json
import os

import numpy as np
import pytest
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TAGS = ["sym5", "sym12", "gen9", "sym40"]


@pytest.mark.parametrize("tag", TAGS)
def test_oracle_cal_loss_matches_reference(golden, orc, tag):
    g = golden("graph")
    A, B, X = g[f"{tag}_A"], g[f"{tag}_B"], g[f"{tag}_X"]
    got = np.array([orc.cal_loss(x, A, B) for x in X])
    assert np.allclose(got, g[f"{tag}_loss"], rtol=1e-13, atol=1e-12)


@pytest.mark.parametrize("tag", TAGS)
def test_mirror_cal_loss_matches_reference(golden, tag):
    import graph
    g = golden("graph")
    A, B, X = g[f"{tag}_A"], g[f"{tag}_B"], g[f"{tag}_X"]
    got = np.array([graph.cal_loss(x, A, B) for x in X])
    assert np.allclose(got, g[f"{tag}_loss"], rtol=1e-12, atol=1e-11)


def test_oracle_enumeration_is_a_maximiser(golden, orc):
    g = golden("graph")
    for tag in ("sym5", "sym12", "gen9"):
        A, B = g[f"{tag}_A"], g[f"{tag}_B"]
        x, obj, n_opt = orc.miqp_bruteforce(A, B, count_optimal=True)
        assert obj == pytest.approx(orc.cal_loss(x, A, B), rel=1e-13)
        assert obj >= g[f"{tag}_loss"].max() - 1e-12            # beats every recorded random assignment
        assert n_opt % 2 == 0                                    # x and 1 - x always tie
        assert orc.cal_loss(1 - x, A, B) == pytest.approx(obj, rel=1e-13)


def test_oracle_recovers_a_planted_orientation(orc):
    """Unfrustrated graph: w = +|w| between patches of equal planted state, -|w| otherwise, invw = -w.
    The planted assignment (and its twin) is the unique optimum and reaches sum |w| * 2."""
    rng = np.random.default_rng(5)
    n = 14
    planted = rng.integers(0, 2, n)
    W = np.triu(rng.uniform(0.1, 1.0, (n, n)), 1) * (rng.random((n, n)) < 0.6)
    W[np.arange(n - 1), np.arange(1, n)] = 0.5                  # keep it connected
    S = np.where(planted[:, None] == planted[None, :], 1.0, -1.0)
    A = (W + W.T) * S
    x, obj, n_opt = orc.miqp_bruteforce(A, -A, count_optimal=True)
    assert n_opt == 2
    assert np.array_equal(x, planted) or np.array_equal(1 - x, planted)
    assert obj == pytest.approx(np.abs(A).sum(), rel=1e-13)


def test_container_matches_reference(golden, orc):
    import graph
    g = golden("graph")
    G = graph.BidGraph()
    G.V = list(range(7))
    for u, v, w, invw in g["container_edges"]:
        G.E.append(graph.BiEdge(int(u), int(v), np.float32(w), np.float32(invw)))
    A, B = G.to_matrix()
    assert A.dtype == np.float64 and np.array_equal(A, g["container_A"]) and np.array_equal(B, g["container_B"])
    A2, B2 = orc.bidgraph_to_matrix(7, [(int(u), int(v), w, iw) for u, v, w, iw in g["container_edges"]])
    assert np.array_equal(A2, A) and np.array_equal(B2, B)
    assert list(G.E[1]) == [1, 2, np.float32(-1.25), np.float32(1.25)]
    G.V = [0, 1, 3]
    with pytest.raises(AssertionError):
        G.to_matrix()


def test_bfs_routes_match_reference(golden, capsys):
    import graph
    g = golden("graph")
    L = graph.LinkedListGraph(int(g["bfs_n"]))
    for u, v, w in g["bfs_edges"]:
        L.add_edge(int(u), int(v), float(w))
    assert np.array_equal(L.get_bfs_route(0), g["bfs_route_0"])
    assert np.array_equal(L.get_bfs_route(33), g["bfs_route_33"])
    assert np.array_equal(L.get_weighted_bfs_route(0), g["bfs_weighted_route_0"])
    assert np.array_equal(L.get_weighted_bfs_route(45), g["bfs_weighted_route_45"])
    assert "bfs warning" in capsys.readouterr().out              # several components, as the reference reports


def test_signatures_match_the_reference():
    import graph
    want = json.load(open(os.path.join(GOLDEN, "graph_signatures.json")))
    for name, params in want.items():
        obj = graph
        for part in name.split("."):
            obj = getattr(obj, part)
        got = [[p.name, None if p.default is inspect.Parameter.empty else repr(p.default)]
               for p in inspect.signature(obj).parameters.values() if p.kind is not inspect.Parameter.VAR_KEYWORD]
        assert got == params, name


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_miqp_has_no_cpu_fallback():
    import dnp_native
    import graph
    A = np.array([[0.0, 1.0], [1.0, 0.0]])
    with pytest.raises(dnp_native.DipoleError):
        graph.MIQP(A, -A)


# ------------------------------------------------------------------ patch graph bookkeeping (util.py:85-108,160-212,418-447)
def _csr(parts):
    parts = [np.asarray(p) for p in parts]
    return np.concatenate(parts), np.concatenate([[0], np.cumsum([len(p) for p in parts])])


@pytest.mark.parametrize("tag", ["hand", "fandisk"])
def test_lzd_partition_and_merge_match_reference(golden, tag):
    """Index work: identical to the reference's own outputs (tests/golden/patchgraph.npz)."""
    import dnp_util
    g = golden("patchgraph")
    pc = torch.from_numpy(g[f"{tag}_pc"])
    n_part, min_patch = int(g[f"{tag}_n_part"]), int(g[f"{tag}_min_patch"])
    idx, ijk = dnp_util._lzd_divide_pc(pc, n_part, min_patch=min_patch)
    assert all(len(c) == 1 and c[0].dtype == torch.int64 for c in idx) and all(len(c) == 1 for c in ijk)
    perm, off = _csr([c[0] for c in idx])
    assert np.array_equal(perm, g[f"{tag}_cell_perm"]) and np.array_equal(off, g[f"{tag}_cell_offsets"])
    assert np.array_equal(torch.stack([c[0] for c in ijk]).numpy(), g[f"{tag}_cell_ijk"])
    m_idx, m_ijk, m_src = dnp_util.lzd_merge_nodes(pc, idx, ijk, min_patch)
    perm, off = _csr(m_idx)
    assert np.array_equal(perm, g[f"{tag}_merged_perm"]) and np.array_equal(off, g[f"{tag}_merged_offsets"])
    assert np.array_equal(torch.stack(m_ijk).numpy(), g[f"{tag}_merged_ijk"])
    assert np.array_equal(np.concatenate([torch.stack(c).numpy() for c in m_src]), g[f"{tag}_source_ijk"])
    assert np.array_equal(np.concatenate([[0], np.cumsum([len(c) for c in m_src])]), g[f"{tag}_source_offsets"])
    # the adjacent patch pairs, in the reference's (i, j) order
    assert np.array_equal(np.array(dnp_util.graph_edges(m_src)), g[f"{tag}_graph_edges"])


def test_util_signatures_match_the_reference():
    import dnp_util
    want = json.load(open(os.path.join(GOLDEN, "util_signatures.json")))
    for name, params in want.items():
        if not hasattr(dnp_util, name):
            continue                                             # xyz2tensor / export_pc: see tests/test_io.py
        got = [[p.name, None if p.default is inspect.Parameter.empty else repr(p.default)]
               for p in inspect.signature(getattr(dnp_util, name)).parameters.values()]
        assert got == params, name
