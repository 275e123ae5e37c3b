"""Drop-in for the dipole hot path of the reference's field_utils.py
(crazyMessi/dipole-normal-prop @ 431bb62), backed by libdipole.so (sm_100a).

Put this directory in front of the reference checkout on PYTHONPATH and
orient_simple.py / orient_pointcloud.py / orient_large.py /
reference_orientation.py / dipole_api.py / graph_dipole.py pick it up through
their `import field_utils` / `from field_utils import *` unchanged: names,
positional order, keyword names, defaults and return conventions are the
reference's.  The greedy propagation control flow
stays in Python/PyTorch as in the reference; every all-pairs evaluation, patch
reduction and sign test is one call into the CUDA library.

There is NO CPU fallback.  CPU tensors are accepted the way the reference
accepts them (result on the input's device), but the arithmetic always runs on
the current CUDA device through the host-buffer entry points of the C ABI.
float64 inputs raise TypeError unless ALLOW_FLOAT64_CAST is set (the kernels
are FP32; SURVEY.md section 7 "dtype").

Known, documented divergences (DESIGN.md "Parity"):
  * `potential`: a coincident (source, target) pair contributes 0; the reference
    zeroes the whole leaf total of that target instead (field_utils.py:53-54), a
    result that depends on its chunking.  Degenerate input only.
  * `recursive` / `max_pts` are accepted and ignored: memory is O(S + M).
  * the NaN/Inf warnings of field_grad (field_utils.py:110-113) are printed only when
    REPORT_NONFINITE is set (one host sync per call); non-finite components are always
    zeroed in the kernel epilogue.
"""
import torch
import numpy as np

try:  # the reference re-exports `util` through `from field_utils import *` (orient_simple.py:5,13)
    import util
except Exception:  # reference checkout (or its open3d dependency) not importable: hot-path helpers only
    import dnp_util as util

import dnp_native as _dn

ALLOW_FLOAT64_CAST = False
# field_utils.py:53-54: a target that coincides with a source makes the reference's LEAF total for that target NaN,
# which it replaces by 0 -- the target loses the contribution of every source that shares a max_pts-leaf with its twin,
# a result that depends on the recursive chunking.  The kernels drop only the coincident PAIR (the documented
# default).  With this switch `potential` reproduces the reference's value on such degenerate inputs: it re-creates
# the leaf decomposition of `recursive` / `max_pts` and subtracts, for every affected target, the partial sums of the
# leaves that hold a twin (extra launches only for affected targets).
REFERENCE_NAN_SEMANTICS = False
# field_utils.py:110-113 prints how many Inf / NaN components field_grad zeroed, at the price of two host
# synchronisations per leaf.  Here the count comes from a device counter of the kernel epilogue; reading it is one
# synchronisation per call, so the warning is opt-in.
REPORT_NONFINITE = False
NATIVE_PATCH_LOOP = True      # greedy patch loops behind one native call (dnp_patch_propagate_f32)

_EPS_DEFAULT = 1e-5
_EPS_POINTS = 1e-6      # field_utils.py:362,380


# ---------------------------------------------------------------------------- plumbing
def _cuda_device():
    _dn.require_cuda()
    return torch.device("cuda", torch.cuda.current_device())


def _check_dtype(t, name):
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a torch.Tensor (got {type(t).__name__})")
    if t.dtype == torch.float32:
        return t, None
    if t.dtype == torch.float64 and ALLOW_FLOAT64_CAST:
        return t.float(), torch.float64
    raise TypeError(f"{name} must be float32 (got {t.dtype}); the B200 kernels compute in FP32 "
                    "(set field_utils.ALLOW_FLOAT64_CAST = True to cast explicitly)")


def _both_cuda(a, b):
    """True if both tensors are CUDA, False if both are CPU (served by the host-buffer ABI calls)."""
    if a.is_cuda != b.is_cuda:
        raise ValueError("sources and means must be on the same device")
    return a.is_cuda


class _Work:
    """A CUDA float32 [N,6] working tensor for the in-place drivers.  It IS the caller's tensor when that is
    already a CUDA float32 tensor with unit column stride (the reference mutates its argument in place:
    orient_simple.py:24 ignores the return value); otherwise a device copy whose normals are written back."""

    def __init__(self, pts, name="pts"):
        t, self.cast_back = _check_dtype(pts, name)
        if t.dim() != 2 or t.shape[1] != 6:
            raise ValueError(f"{name} must be [N,6] (got {tuple(pts.shape)})")
        self.orig = pts
        if t is pts and pts.is_cuda and (pts.shape[0] <= 1 or (pts.stride(1) == 1 and pts.stride(0) >= 6)):
            self.t = pts
            self.alias = True
        else:
            dev = pts.device if pts.is_cuda else _cuda_device()
            self.t = t.detach().to(dev, copy=True).contiguous()
            self.alias = False

    def finish(self):
        if not self.alias:
            self.orig[:, 3:] = self.t[:, 3:].to(device=self.orig.device, dtype=self.orig.dtype)


def _clear(mask, j):
    """mask[j] = 0 for a Python int j without the host-to-device scalar copy `mask[j] = 0` makes."""
    mask.narrow(0, int(j), 1).zero_()


def _index_list(idx_list, dev):
    out = []
    for idx in idx_list:
        if torch.is_tensor(idx) and idx.dtype == torch.int64 and idx.device == dev and idx.dim() == 1:
            out.append(idx)                       # the common case, without the per-tensor dispatch overhead
            continue
        idx = torch.as_tensor(idx)
        if idx.dtype == torch.bool:
            idx = torch.nonzero(idx).flatten()
        out.append(idx.to(device=dev, dtype=torch.int64).flatten())
    return out


def _csr(idx_list, dev):
    """(perm, offsets) of a list of index tensors: perm[offsets[p]:offsets[p+1]] are the rows of patch p."""
    sizes = [int(i.numel()) for i in idx_list]
    offsets = torch.zeros(len(idx_list) + 1, dtype=torch.int64)
    if sizes:
        offsets[1:] = torch.tensor(sizes, dtype=torch.int64).cumsum(0)
    perm = torch.cat(idx_list) if idx_list else torch.zeros(0, dtype=torch.int64, device=dev)
    return perm.contiguous(), offsets.to(dev)


def _flattest_patch(pts, perm, offsets):
    """argmin_p |smallest eigenvalue of cov(xyz of patch p)|, first on ties (field_utils.py:230-233,303-306).
    The covariances come from the device in float64 (fixed order) and are rounded to float32 before a float32
    `eigh`, the precision util.pca_eigen_values works in (util.py:495-500): on clouds with several exactly planar
    patches lambda_0 is rounding noise, and its magnitude -- hence the arg-min -- depends on that precision.
    Known divergence (DESIGN.md "Parity"): the float32 covariance itself is accumulated in another order than
    `temp.T @ temp`, so between exactly planar patches the start patch can still differ from the reference's."""
    cov = _dn.patch_covariance(pts, perm, offsets).to(torch.float32).cpu()
    lam0 = torch.linalg.eigvalsh(cov)[:, 0].numpy()
    return int(np.argmin(np.abs(lam0)))


# ---------------------------------------------------------------------------- a1-a3: leaf functions
def measure_mean_potential(pc: torch.Tensor):
    """field_utils.py:7-9: mean dipole potential over the 10x10x10 lattice of util.gen_grid (util.py:26-36).
    The lattice is generated inside the kernel.  Returns a 0-d tensor on pc's device."""
    t, cast = _check_dtype(pc, "pc")
    if t.is_cuda:
        out = _dn.lattice_mean_potential(t)
    else:
        out = _dn.lattice_mean_potential(t.to(_cuda_device())).cpu()
    return out.to(cast) if cast else out


def potential(sources, means, eps=1e-5, recursive=True, max_pts=15000):
    """field_utils.py:12-55.  phi[m] = sum_s n_s . (p_s - q_m) / |p_s - q_m|^3.
    `eps` is unused by the reference as well; `recursive` / `max_pts` are ignored (no chunking needed)."""
    s, cast = _check_dtype(sources, "sources")
    m, cast_m = _check_dtype(means, "means")
    if REFERENCE_NAN_SEMANTICS:
        dev = s.device if s.is_cuda else _cuda_device()
        out = _potential_with_leaf_laundering(s.to(dev), m.to(dev), recursive, max_pts).to(m.device)
    elif _both_cuda(s, m):
        out = _dn.potential(s, m)
    else:
        out = _dn.potential_host(s, m)
    return out.to(cast) if cast else out


def _leaves(S, M, recursive, max_pts):
    """(source range, target range) of every leaf of the reference's recursion (field_utils.py:23-44)."""
    todo, out = [(0, S, 0, M)], []
    while todo:
        sa, sb, ta, tb = todo.pop()
        ns, nt = sb - sa, tb - ta
        split_src = recursive and ns > max_pts and (nt <= max_pts or ns > nt)
        split_tgt = recursive and not split_src and nt > max_pts
        if split_src:
            mid = sa + int(ns / 2)
            todo += [(sa, mid, ta, tb), (mid, sb, ta, tb)]
        elif split_tgt:
            mid = ta + int(nt / 2)
            todo += [(sa, sb, ta, mid), (sa, sb, mid, tb)]
        else:
            out.append((sa, sb, ta, tb))
    return out


def _potential_with_leaf_laundering(s, m, recursive, max_pts):
    """REFERENCE_NAN_SEMANTICS: the pair-dropping potential minus, per target, the partial sums of the leaves in
    which one of its coincident sources lies (those leaf totals are NaN -> 0 in the reference)."""
    phi = _dn.potential(s, m)
    S, M = s.shape[0], m.shape[0]
    if S == 0 or M == 0:
        return phi
    # exact coordinate coincidences (+0.0 == -0.0, NaN never equal): rows grouped by value
    xyz = torch.cat([s[:, :3], m[:, :3]], 0) + 0.0
    finite = ~torch.isnan(xyz).any(dim=1)
    _, inv = torch.unique(xyz, dim=0, return_inverse=True)
    src_id, tgt_id = inv[:S], inv[S:]
    present = torch.zeros(int(inv.max()) + 1, dtype=torch.bool, device=s.device)
    present[src_id[finite[:S]]] = True
    hit_t = torch.nonzero(present[tgt_id] & finite[S:]).flatten()
    if hit_t.numel() == 0:
        return phi
    # (target, source) twin pairs: sort the sources by group id once, then look every hit target's group up
    order = torch.argsort(src_id, stable=True)
    sorted_ids = src_id[order]
    lo = torch.searchsorted(sorted_ids, tgt_id[hit_t])
    hi = torch.searchsorted(sorted_ids, tgt_id[hit_t], right=True)
    counts = hi - lo
    t_rep = torch.repeat_interleave(hit_t, counts)
    pos = torch.arange(int(counts.sum()), device=s.device) - torch.repeat_interleave(torch.cumsum(counts, 0) - counts, counts)
    s_rep = order[torch.repeat_interleave(lo, counts) + pos]
    # affected targets are rebuilt leaf by leaf, as the reference sums them: a leaf holding a twin contributes 0
    phi[hit_t] = 0
    for sa, sb, ta, tb in _leaves(S, M, bool(recursive), int(max_pts)):
        rows = hit_t[(hit_t >= ta) & (hit_t < tb)]
        if rows.numel() == 0:
            continue
        twin_here = (t_rep >= ta) & (t_rep < tb) & (s_rep >= sa) & (s_rep < sb)
        dirty = torch.zeros(M, dtype=torch.bool, device=s.device)
        dirty[t_rep[twin_here]] = True
        rows = rows[~dirty[rows]]
        if rows.numel():
            phi[rows] += _dn.potential(s[sa:sb], m[rows].contiguous())
    return phi


def field_grad(sources, means, eps=1e-5, recursive=True, max_pts=15000):
    """field_utils.py:61-116.  E[m] = -sum_s (3 (n_s.Rh) Rh - n_s) / (|R|^3 + eps), R = p_s - q_m;
    coincident pairs contribute exactly 0 (:98-106); non-finite components are zeroed (:114-115)."""
    s, cast = _check_dtype(sources, "sources")
    m, cast_m = _check_dtype(means, "means")
    if _both_cuda(s, m):
        if REPORT_NONFINITE:
            counter = torch.zeros(1, dtype=torch.int32, device=m.device)
            out = _dn.field_grad(s, m, eps=eps, nonfinite=counter)
            bad = int(counter.item())
            if bad:
                print("warning: %d inf / nan in field_grad" % bad)            # field_utils.py:110-113
        else:
            out = _dn.field_grad(s, m, eps=eps)
    else:
        out = _dn.field_grad_host(s, m, eps=eps)
    return out.to(cast) if cast else out


# ---------------------------------------------------------------------------- a4: patch-pair flip scores
def _edge_interaction(S, T):
    s, _ = _check_dtype(S, "sources")
    t, _ = _check_dtype(T, "means")
    if s.shape[1] < 6 or t.shape[1] < 6:
        raise ValueError("both patches need 6 columns")
    if not _both_cuda(s, t):
        dev = _cuda_device()
        s, t = s.to(dev), t.to(dev)
    return _dn.edge_interaction(s, t, eps=_EPS_DEFAULT)


def field_edge_calculator(sources, means, if_save=False):
    """field_utils.py:145-160: w = (2 sum_t n_t.E_S(t)) / |S| * |T| (precedence as written at :155);
    returns (numpy 0-d float32 array, numpy float32) = (w, -w)."""
    st_interaction = _edge_interaction(sources, means)
    w = (st_interaction * 2) / sources.shape[0] * means.shape[0]
    w = w.detach().cpu().numpy()
    invw = w * -1
    return w, invw


def field_edge_calculator_bool(sources, means, if_save=False):
    """field_utils.py:129-134."""
    w, invw = field_edge_calculator(sources, means, if_save)
    if w > 0:
        return 1, -1
    else:
        return -1, 1


def field_edge_calculator_count(sources, means, if_save=False):
    """field_utils.py:137-143 (the score is evaluated, then replaced by |S||T|)."""
    w, invw = field_edge_calculator(sources, means, if_save)
    w = sources.shape[0] * means.shape[0]
    if w > 0:
        return w, -w
    else:
        return -w, w


def field_edge_calculator_batched(pc, patches, edges):
    """All edges of a patch graph in ONE launch (K3; no reference counterpart -- the reference calls
    field_edge_calculator once per edge from Python threads, util.py:199-211,308-327).
    pc [N,6]; patches: list of index tensors; edges: [nE,2] (source patch, target patch).
    Returns (w, invw) float32 numpy arrays laid out like BiEdge.w / BiEdge.invw (graph.py:121-129)."""
    work, _ = _check_dtype(pc, "pc")
    dev = work.device if work.is_cuda else _cuda_device()
    work = work.to(dev)
    idx = _index_list(patches, dev)
    perm, offsets = _csr(idx, dev)
    edges = torch.as_tensor(edges, dtype=torch.int64).reshape(-1, 2).to(dev).contiguous()
    inter = _dn.edge_interactions_batched(work, perm, offsets, edges, eps=_EPS_DEFAULT)
    sizes = (offsets[1:] - offsets[:-1]).to(torch.float32)
    w = (inter * 2) / sizes[edges[:, 0]] * sizes[edges[:, 1]]
    w = w.cpu().numpy()
    return w, w * -1


def self_interaction(nxyz, eps=1e-5):
    """field_utils.py:163-171: score between a random half of the patch and the other half."""
    assert nxyz.shape[1] == 6
    num = nxyz.shape[0]
    mask = torch.ones(num, dtype=torch.bool)
    mask[torch.randperm(nxyz.shape[0])[:int(num / 2)]] = False
    mask = mask.to(nxyz.device)
    w, _ = field_edge_calculator(nxyz[mask], nxyz[~mask])
    return w


def self_interaction_all(nxyz, eps=1e-5):
    """field_utils.py:174-177."""
    assert nxyz.shape[1] == 6
    w, _ = field_edge_calculator(nxyz, nxyz)
    return w


def random_self_interaction(nxyz, eps=1e-5):
    """field_utils.py:179-186: self score after flipping a random half of the normals."""
    assert nxyz.shape[1] == 6
    rand_flip = np.zeros(nxyz.shape[0], dtype=bool)
    rand_flip[np.random.permutation(nxyz.shape[0])[:int(nxyz.shape[0] / 2)]] = True
    rand_n = nxyz.clone()
    rand_n[torch.from_numpy(rand_flip).to(nxyz.device), 3:] *= -1
    w, _ = field_edge_calculator(rand_n, rand_n)
    return w


# ---------------------------------------------------------------------------- a5: orientation transfer
def reference_field(pc1, pc2):
    """field_utils.py:188-201.  3-column target: returns a NEW [M,6] tensor with normals E/|E| (rows with
    |E| = 0 keep the raw E).  6-column target: flips pc2's normals IN PLACE where E.n < 0 and returns pc2."""
    with torch.no_grad():
        src, _ = _check_dtype(pc1, "pc1")
        tgt, cast = _check_dtype(pc2, "pc2")
        if tgt.dim() != 2 or tgt.shape[1] not in (3, 6):
            raise ValueError("pc2 must have 3 or 6 columns")
        dev = tgt.device if tgt.is_cuda else _cuda_device()
        src_d = src.to(dev)
        if tgt.shape[1] == 3:
            out = torch.empty(tgt.shape[0], 6, dtype=torch.float32, device=dev)
            out[:, :3] = tgt.to(dev)
            _dn.field_apply(src_d, out, _EPS_DEFAULT, _dn.EPI_NORMALIZE, out=out[:, 3:])
            out = out.to(pc2.device)
            return out.to(cast) if cast else out
        work = _Work(pc2, "pc2")
        _dn.field_apply(src_d.to(work.t.device), work.t, _EPS_DEFAULT, _dn.EPI_SIGN_GE, normals=work.t[:, 3:])
        work.finish()
        return pc2


# ---------------------------------------------------------------------------- the "xie" interaction
def _xie_inputs(source, target, knn_mask, need_target_normals):
    s, cast = _check_dtype(source, "source")
    t, _ = _check_dtype(target, "target")
    if s.dim() != 2 or s.shape[1] < 6:
        raise ValueError("source must be [N, 6]")
    if t.dim() != 2 or t.shape[1] < (6 if need_target_normals else 3):
        raise ValueError("target must be [M, 6]" if need_target_normals else "target must be [M, >=3]")
    on_cuda = _both_cuda(s, t)
    dev = t.device if on_cuda else _cuda_device()
    s_d, t_d = s.to(dev), t.to(dev)
    radius2 = _dn.knn_radius2(s_d, t_d, knn_mask) if knn_mask is not None and knn_mask > 0 else None
    return s_d, t_d, radius2, cast, source.device


def xie_field(source: torch.Tensor, target: torch.Tensor, eps, max_pts=5000, knn_mask=-1, C=3):
    """field_utils.py:431-469.  [M, N, 3]: out[t, s] = (n_s - C (n_s.Rh) Rh) / |R|^3; a coincident pair keeps n_s.
    `eps` is unused by the reference as well; `max_pts` (its chunking) is ignored.  Divergence: the reference's
    recursion drops `knn_mask` and `C` for inputs above max_pts^2 pairs (:436-441 pass neither on); here they
    apply at every size."""
    with torch.no_grad():
        s, t, radius2, cast, dev = _xie_inputs(source, target, knn_mask, False)
        out = _dn.xie_matrix(s, t, C=C, radius2=radius2, field=True).to(dev)
        return out.to(cast) if cast else out


def xie_intersaction(source: torch.Tensor, target: torch.Tensor, eps, knn_mask, C):
    """field_utils.py:509-519.  [M, N]: inter[t, s] = xie_field[t, s] . n_t, NaN / Inf -> 0 (no warning print)."""
    with torch.no_grad():
        s, t, radius2, cast, dev = _xie_inputs(source, target, knn_mask, True)
        out = _dn.xie_matrix(s, t, C=C, radius2=radius2, field=False).to(dev)
        return out.to(cast) if cast else out


def xie_propagation_points_in_order(pts: torch.Tensor, eps, order, diffuse=False, verbose=False, points_weight=None,
                                    knn_mask=-1, C=3):
    """field_utils.py:569-605.  order [T, N]: T visiting orders; returns the bool tensor [T, N] `interactions < 0`
    (flip status of every point in every chain).  The [N, N] interaction matrix is never formed: one cluster per
    chain evaluates the visited point's row on the fly (dnp_xie_propagate_in_order_f32).  `points_weight` has no
    effect in the reference either (:584-585 multiplies by ones); `eps` and `verbose` are unused."""
    with torch.no_grad():
        p, _ = _check_dtype(pts, "pts")
        if p.dim() != 2 or p.shape[1] != 6:
            raise ValueError("pts must be [N, 6]")
        dev = p.device if p.is_cuda else _cuda_device()
        p_d = p.to(dev)
        radius2 = _dn.knn_radius2(p_d, p_d, knn_mask) if knn_mask is not None and knn_mask > 0 else None
        order_t = torch.as_tensor(np.asarray(order) if not isinstance(order, torch.Tensor) else order).long()
        flips, _ = _dn.xie_propagate_in_order(p_d, order_t, C=C, radius2=radius2, diffuse=diffuse)
        return flips.to(pts.device)


# ---------------------------------------------------------------------------- flip vote
def cal_loss(x, A, B):
    """field_utils.py:606-615 (same sum as graph.cal_loss)."""
    import graph
    return graph.cal_loss(x, A, B)


def MIQP(A, B):
    """field_utils.py:618-643.  NOT graph.MIQP: this one sets the objective to `-cal_loss` (:628-629), i.e. it
    returns the x in {0,1}^n MINIMISING cal_loss(x, A, B), and prints Gurobi's objective value (-cal_loss).
    Solved on the GPU by dnp_flip_solve_f64 on (-A, -B)."""
    import graph
    x, obj, _, _ = graph._solve(A, B, False)
    print('Obj: %g' % -obj)
    return x


def xie_propagation_points_onbfstree(pts: torch.Tensor, eps, diffuse=False, starting_point=0, verbose=False, k=10,
                                     treshold=0.1, times=1, use_pw=False, knn_mask=-1, C=3):
    """field_utils.py:654-709.  `times` BFS orders over the thresholded kNN graph, one propagation chain per order
    (one launch, a cluster per chain), the chains aligned by the flip vote (MIQP on the pairwise disagreement
    counts), then a per-point majority vote flips pts[:, 3:] in place.  Returns the per-point flip mask
    `cnts > times / 2` (bool [N], on pts' device) like the reference (:709)."""
    import graph
    assert times % 2 == 1 and times > 0
    with torch.no_grad():
        starting_points = [starting_point]
        np.random.seed(0)                                    # :660, the reference reseeds numpy here
        while len(np.unique(starting_points)) < times:
            t = np.random.randint(0, pts.shape[0])
            if t not in starting_points:
                starting_points.append(t)
        xyz = pts[:, :3].cpu().numpy()
        G, mean_k_dist = graph.getEMSTfromPC(xyz, k, treshold)
        point_weight = mean_k_dist if use_pw else None
        orders = np.zeros([times, len(pts)], dtype=int)
        for i in range(times):
            orders[i] = G.get_bfs_route(starting_points[i])
        status_all = xie_propagation_points_in_order(pts.clone(), eps, orders, diffuse, verbose=False,
                                                     points_weight=point_weight, knn_mask=knn_mask, C=C).T   # [N, times]
        # pairwise disagreement / agreement counts of the chains (:682-686,696-700) as one integer mat-mul
        f = status_all.to(torch.float64)
        n = f.shape[0]
        diff = f.T @ (1.0 - f)
        A = diff + diff.T
        B = n - A
        status = MIQP(A.cpu().numpy(), B.cpu().numpy())
        status = torch.as_tensor(status != 0, device=status_all.device)
        aligned = status_all ^ status[None, :]
        cnts = aligned.int().sum(dim=1)
        flipped = (cnts > times / 2).to(pts.device)
        pts[flipped, 3:] *= -1
    return flipped


def _reps_loop_per_call(pts, rep_idx, rest_idx, perm, offsets, start, E, diffuse):
    """The loop of field_utils.py:239-265 with one library call per kernel (the first version of this driver;
    kept as the reference for the native loop, tests/test_gpu_drivers.py)."""
    dev = pts.device
    N, P = pts.shape[0], len(rep_idx)
    settled = torch.ones(N, dtype=torch.uint8, device=dev)     # NOT non_oriented_pts_mask (:223-227)
    settled.index_fill_(0, perm, 0)
    scratch_mask = torch.zeros(N, dtype=torch.uint8, device=dev)
    active = torch.ones(P, dtype=torch.uint8, device=dev)
    patch = rep_idx[start]
    _clear(active, start)
    settled.index_fill_(0, patch, 1)
    _dn.field_apply(pts, pts, _EPS_DEFAULT, _dn.EPI_STORE, out=E, src_idx=patch,          # :239
                    tgt_idx=perm, tgt_skip=settled)
    for _ in range(P - 1):
        _, best = _dn.patch_interactions(E, pts, perm, offsets, active)                   # :244-247
        j, value = _dn.best_to_host(best)
        patch = rep_idx[j]
        if value < 0:                                                                     # :250-252
            pts[patch, 3:] *= -1
            pts[rest_idx[j], 3:] *= -1
        _clear(active, j)
        settled.index_fill_(0, patch, 1)
        if diffuse:                                                                       # :256-261
            scratch_mask.index_fill_(0, patch, 1)
            _dn.field_apply(pts, pts, _EPS_DEFAULT, _dn.EPI_ACCUM, out=E, src_idx=patch,
                            tgt_idx=perm, tgt_skip=scratch_mask)
            scratch_mask.index_fill_(0, patch, 0)
        else:                                                                             # :264-265
            _dn.field_apply(pts, pts, _EPS_DEFAULT, _dn.EPI_ACCUM, out=E, src_idx=patch,
                            tgt_idx=perm, tgt_skip=settled)


def _patch_loop_per_call(p, idx, perm, offsets, start, E, diffuse):
    """The loop of field_utils.py:311-335 with one library call per kernel (see _reps_loop_per_call)."""
    dev = p.device
    N, P = p.shape[0], len(idx)
    oriented = torch.zeros(N, dtype=torch.uint8, device=dev)      # pts_mask (:296)
    scratch_mask = torch.zeros(N, dtype=torch.uint8, device=dev)
    active = torch.ones(P, dtype=torch.uint8, device=dev)
    patch = idx[start]
    _clear(active, start)
    oriented.index_fill_(0, patch, 1)
    _dn.field_apply(p, p, _EPS_DEFAULT, _dn.EPI_STORE, out=E, src_idx=patch, tgt_skip=oriented)   # :311
    for _ in range(P - 1):
        _, best = _dn.patch_interactions(E, p, perm, offsets, active)                     # :316-319
        j, value = _dn.best_to_host(best)
        patch = idx[j]
        if value < 0:                                                                     # :322-323
            p[patch, 3:] *= -1
        _clear(active, j)
        oriented.index_fill_(0, patch, 1)
        if diffuse:                                                                       # :326-331
            scratch_mask.index_fill_(0, patch, 1)
            _dn.field_apply(p, p, _EPS_DEFAULT, _dn.EPI_ACCUM, out=E, src_idx=patch, tgt_skip=scratch_mask)
            scratch_mask.index_fill_(0, patch, 0)
        else:                                                                             # :333-335
            _dn.field_apply(p, p, _EPS_DEFAULT, _dn.EPI_ACCUM, out=E, src_idx=patch, tgt_skip=oriented)


# ---------------------------------------------------------------------------- a8: patch propagation on representatives
def strongest_field_propagation_reps(input_pc, reps, diffuse=False, weights=None):
    """field_utils.py:207-282.  reps = [(rep_idx, rest_idx)] per patch.  Mutates input_pc[:, 3:]; returns None."""
    input_pc = input_pc.detach()
    with torch.no_grad():
        work = _Work(input_pc, "input_pc")
        pts = work.t
        dev = pts.device
        N = pts.shape[0]
        if weights is not None:
            weights = weights.to(device=dev, dtype=torch.float32).clamp(0.1, 1)          # :212-215
            pts[:, 3:] = pts[:, 3:] * weights[:, None]

        rep_idx = _index_list([rep for rep, rest in reps], dev)
        rest_idx = _index_list([rest for rep, rest in reps], dev)
        P = len(rep_idx)
        perm, offsets = _csr(rep_idx, dev)            # perm = every representative row
        rest_perm, rest_offsets = _csr(rest_idx, dev)
        E = torch.zeros(N, 3, dtype=torch.float32, device=dev)
        oriented = torch.zeros(N, dtype=torch.uint8, device=dev)
        oriented.index_fill_(0, perm, 1)              # after the loop every representative is oriented

        if P > 0:
            start = _flattest_patch(pts, perm, offsets)                                   # :230-233
            if NATIVE_PATCH_LOOP:
                # :239-265 in one native call: interactions, arg-max, flips (reps + the rest of the patch), field
                _dn.patch_propagate(pts, perm, offsets, start, E, diffuse=diffuse, eps=_EPS_DEFAULT,
                                    rest_perm=rest_perm, rest_offsets=rest_offsets, tgt_idx=perm)
            else:
                _reps_loop_per_call(pts, rep_idx, rest_idx, perm, offsets, start, E, diffuse)

        if diffuse:                                                                       # :267-271
            _dn.field_sign(E, pts, idx=perm, strict=True)

        rest_rows = torch.nonzero(oriented == 0).flatten()                                # :273-276
        if rest_rows.numel() > 0:
            src_rows = torch.nonzero(oriented).flatten()
            _dn.field_apply(pts, pts, _EPS_DEFAULT, _dn.EPI_SIGN_GT, normals=pts[:, 3:],
                            src_idx=src_rows, tgt_idx=rest_rows)

        if weights is not None:                                                           # :280-282
            pts[:, 3:] = pts[:, 3:] / weights[:, None]
        work.finish()


# ---------------------------------------------------------------------------- a7: patch propagation on all points
def strongest_field_propagation(pts, patches, all_patches, diffuse=False, weights=None):
    """field_utils.py:286-348.  patches = [(i, idx)] network-kept, all_patches = [idx].  Mutates pts[:, 3:]."""
    with torch.no_grad():
        work = _Work(pts, "pts")
        p = work.t
        dev = p.device
        N = p.shape[0]
        if weights is not None:
            weights = weights.to(device=dev, dtype=torch.float32).clamp(0.1, 1)          # :288-291
            p[:, 3:] = p[:, 3:] * weights[:, None]

        idx = _index_list(all_patches, dev)
        P = len(idx)
        perm, offsets = _csr(idx, dev)
        E = torch.zeros(N, 3, dtype=torch.float32, device=dev)
        if P > 0:
            start = _flattest_patch(p, perm, offsets)                                     # :303-306
            if NATIVE_PATCH_LOOP:
                _dn.patch_propagate(p, perm, offsets, start, E, diffuse=diffuse, eps=_EPS_DEFAULT)   # :311-335
            else:
                _patch_loop_per_call(p, idx, perm, offsets, start, E, diffuse)

        if diffuse:                                                                       # :337-342
            kept = _index_list([entry[1] for entry in patches], dev)
            if kept:
                _dn.field_sign(E, p, idx=torch.cat(kept), strict=True)

        if weights is not None:                                                           # :346-348
            p[:, 3:] = p[:, 3:] / weights[:, None]
        work.finish()


# ---------------------------------------------------------------------------- a6: per-point propagation
def strongest_field_propagation_points(pts: torch.Tensor, diffuse=False, starting_point=0, verbose=False):
    """field_utils.py:353-388.  The N-1 greedy steps run inside one persistent cluster kernel
    (dnp_points_propagate_f32) in the reference's float32 evaluation order.  Mutates pts[:, 3:], returns pts.
    `verbose` only opened an interactive viewer in the reference (:366-367) and is ignored."""
    with torch.no_grad():
        work = _Work(pts, "pts")
        _dn.points_propagate(work.t, diffuse=diffuse, starting_point=int(starting_point), eps=_EPS_POINTS)
        work.finish()
    return pts
