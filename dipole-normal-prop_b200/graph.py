"""Drop-in for the part of the reference's graph.py that sits on the flip-score path
(SURVEY.md section 8a row a11 and section 8f row 2):

  BiEdge / BidGraph      graph.py:121-158   score container filled by field_edge_calculator
  cal_loss               graph.py:431-442   objective of the patch flip solve
  MIQP                   graph.py:444-470   the solve itself: Gurobi in the reference, here the sm_100a
                                            kernels of csrc/dnp_flip.cu behind dnp_flip_solve_f64
  LinkedListGraph        graph.py:257-318   adjacency lists + the BFS routes the "xie" propagation walks
  getEMSTfromPC          graph.py:355-367   thresholded kNN graph (sklearn KD-tree, as in the reference)

Not provided: the open3d mesh/visualisation helpers, GraphPC (ground-truth metrics) and the KD-tree timing
toys -- none of them is on the path.  There is no CPU fallback: MIQP raises without a CUDA device.
"""
import heapq

import numpy as np
import torch

import dnp_native


class BiEdge:
    """graph.py:121-129: edge u-v with the score `w` for "same orientation" and `invw` for "one of them flipped"."""

    def __init__(self, u, v, w, invw):
        self.u = u
        self.v = v
        self.w = w
        self.invw = invw

    def __iter__(self):
        return iter([self.u, self.v, self.w, self.invw])


class BidGraph:
    """graph.py:131-158."""

    def __init__(self):
        self.V = []  # node ids
        self.E = []  # BiEdge list

    def to_matrix(self):
        """Dense symmetric A (w) and B (invw), float64 [n, n]; asserts that the ids are 0..n-1 (graph.py:136-152)."""
        ids = set(self.V)
        for i in range(len(ids)):
            if i not in ids:
                print("Error! the graph's vertex id is not continuous")
                assert i in ids
        n = len(self.V)
        A = np.zeros((n, n))
        B = np.zeros((n, n))
        for e in self.E:
            A[e.u][e.v] = A[e.v][e.u] = e.w
            B[e.u][e.v] = B[e.v][e.u] = e.invw
        return A, B

    def add_edge(self, u, v, calculator):
        self.V.append(u)
        self.V.append(v)
        w, invw = calculator(u, v)
        self.E.append(BiEdge(u, v, w, invw))
        return self


def cal_loss(x, A, B):
    """graph.py:431-442.  sum_ij A_ij (1 - (x_i - x_j)^2) + B_ij (x_i - x_j)^2 (host arithmetic: this is the
    reporting function; the solver evaluates the same sum on the device)."""
    n = len(x)
    assert A.shape == (n, n)
    assert B.shape == (n, n)
    x = np.asarray(x, dtype=np.float64)
    d2 = (x[:, None] - x[None, :]) ** 2
    return float((np.asarray(A, dtype=np.float64) * (1.0 - d2) + np.asarray(B, dtype=np.float64) * d2).sum())


def _solve(A, B, maximise, **kw):
    assert A.shape == B.shape
    assert A.shape[0] == A.shape[1]
    dnp_native.require_cuda()
    dev = torch.device("cuda", torch.cuda.current_device())
    sign = 1.0 if maximise else -1.0
    Ad = torch.as_tensor(np.asarray(A, dtype=np.float64)).to(dev) * sign
    Bd = torch.as_tensor(np.asarray(B, dtype=np.float64)).to(dev) * sign
    x, obj, status = dnp_native.flip_solve(Ad, Bd, **kw)
    obj = obj.cpu().numpy() * sign
    return x.cpu().numpy().astype(np.float64), float(obj[0]), float(obj[1]), int(status.item())


LAST_STATUS = {}     # informational: objective / bound / status of the calling thread's last MIQP


def MIQP(A, B, **solver_options):
    """graph.py:444-470: the x in {0,1}^n maximising cal_loss(x, A, B), as a float64 array; prints the objective
    like the reference.  n <= 36: exhaustive, i.e. optimal up to float32 rounding of the energies (the 2^(n-1) walk runs on float32
    couplings; the best 32 candidates are re-evaluated in double).  Larger n: annealing + descent on the GPU; the returned
    point is optimal whenever it reaches the bound sum_pairs max(A, B) (always the case for an unfrustrated
    patch graph), otherwise the best found (`graph.LAST_STATUS`).  Of the two globally flipped twins, which
    have the same objective, the one flipping fewer patches is returned."""
    x, obj, bound, status = _solve(A, B, True, **solver_options)
    LAST_STATUS.update(objective=obj, bound=bound, status=status)
    print('Obj: %g' % obj)
    return x


class LinkedListGraph:
    """graph.py:257-318.  Adjacency sets whose elements hash and compare by end vertex only, so a second edge
    to the same vertex is dropped and iteration follows CPython's small-int set order -- both of which the BFS
    routes of the reference depend on."""

    class halfEdge:
        def __init__(self, v, w):
            self.v = v
            self.w = w

        def __eq__(self, other):
            return self.v == other.v

        def __hash__(self):
            return self.v

        def __lt__(self, other):
            return self.w < other.w

    def __init__(self, node_num):
        self.node_num = node_num
        self.edges = [set() for _ in range(node_num)]

    def add_edge(self, u, v, w):
        self.edges[u].add(LinkedListGraph.halfEdge(v, w))

    def _route(self, start, weighted):
        order = []
        seen = np.zeros(self.node_num, dtype=bool)
        seen[start] = True
        units = 1
        if weighted:
            frontier = [(0, start)]
            push = lambda w, v: heapq.heappush(frontier, (w, v))      # noqa: E731
            pop = lambda: heapq.heappop(frontier)[1]                  # noqa: E731
        else:
            frontier = [start]
            push = lambda w, v: frontier.append(v)                    # noqa: E731
            pop = lambda: frontier.pop(0)                             # noqa: E731
        while frontier:
            u = pop()
            order.append(u)
            for e in self.edges[u]:
                if not seen[e.v]:
                    seen[e.v] = True
                    push(e.w, e.v)
            if not frontier:
                if seen.all():
                    break
                nxt = np.where(~seen)[0][0]          # next component: lowest unvisited vertex
                push(0, nxt)
                seen[nxt] = True
                units += 1
        if units != 1:
            print("bfs warning::unit= ", units)
        return order

    def get_bfs_route(self, start):
        """graph.py:281-304: BFS order from `start`, unweighted."""
        return self._route(start, False)

    def get_weighted_bfs_route(self, start):
        """graph.py:306-331: best-first order by the weight of the discovering edge."""
        return self._route(start, True)


def getEMSTfromPC(xyz: np.ndarray, k: int = 10, threshold: float = 0.1):
    """graph.py:355-367: kNN graph with edges shorter than `threshold`; returns (LinkedListGraph, mean kNN
    distance per point).  The KD-tree query is sklearn's, as in the reference (host plumbing)."""
    from sklearn.neighbors import KDTree
    tree = KDTree(xyz)
    n = len(xyz)
    G = LinkedListGraph(n)
    distance, idx = tree.query(xyz, k)
    for i in range(n):
        for j in range(k):
            if idx[i, j] != i and distance[i, j] < threshold:
                G.add_edge(i, int(idx[i, j]), distance[i, j])
    return G, np.mean(distance, axis=1)
