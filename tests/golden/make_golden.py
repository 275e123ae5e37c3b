#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ from the UNMODIFIED reference.

Run in the build container only (needs /root/reference, which does not exist on
the GPU box):

    python tests/golden/make_golden.py            # all fixtures
    python tests/golden/make_golden.py leaf edge  # a subset

The reference (crazyMessi/dipole-normal-prop @ 431bb62) has no tests and no
golden vectors of its own, so parity is pinned on outputs of the
reference itself run here on its CPU torch path.  The reference module is
imported read-only; the only shims are OUTSIDE the hot-path math:

  * `open3d`, `pymeshlab`, `gurobipy` are absent -> empty stub modules so that
    `import field_utils` succeeds (util.py:7, util.py:533, field_utils.py:619).
  * `torch.symeig` was removed from torch -> `torch.linalg.eigh` (same order),
    needed by util.pca_eigen_values (util.py:495-500).
  * `Tensor.cuda` -> identity: strongest_field_propagation_points calls
    `pts.cuda()` unconditionally (field_utils.py:355); the oracle is the CPU path.
  * `util.divide_pc` is broken at HEAD (util.py:338-341); patches come from
    `util.merge_nodes(pc, *util._divide_pc(...), min_patch)[0]`, the upstream
    composition.
  * The PointCNN stage (out of scope, needs torch_geometric) is replaced by
    seeded random per-patch flips and seeded per-point weights.

To record the greedy decision sequence without touching the drivers, the
module attribute `field_utils.field_grad` is wrapped by a recorder that stores
the sources of every call and then forwards to the original function.
"""
import os
import sys
import time
import types

import numpy as np
import torch

REF = os.environ.get("DNP_REFERENCE", "/root/reference")
OUT = os.path.dirname(os.path.abspath(__file__))


def import_reference():
    for name in ("open3d", "pymeshlab", "gurobipy"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    # torch.symeig survives only as a stub that raises; replace it unconditionally
    torch.symeig = lambda A, eigenvectors=True, upper=True: torch.linalg.eigh(A)
    torch.Tensor.cuda = lambda self, *a, **k: self
    sys.path.insert(0, REF)
    cwd = os.getcwd()
    os.chdir(REF)
    try:
        import field_utils as ref_fu  # noqa
        import util as ref_util  # noqa
    finally:
        os.chdir(cwd)
    return ref_fu, ref_util


def rand_cloud(gen, n, cols=6, scale=1.0):
    xyz = (torch.rand(n, 3, generator=gen) - 0.5) * scale
    if cols == 3:
        return xyz.float()
    nrm = torch.randn(n, 3, generator=gen)
    nrm = nrm / nrm.norm(dim=-1, keepdim=True)
    return torch.cat([xyz, nrm], dim=1).float()


def sphere_cloud(gen, n, r=0.45, noise=0.002, centre=(0.0, 0.0, 0.0)):
    d = torch.randn(n, 3, generator=gen)
    d = d / d.norm(dim=-1, keepdim=True)
    xyz = d * r + torch.randn(n, 3, generator=gen) * noise + torch.tensor(centre)
    return torch.cat([xyz, d], dim=1).float()


def load_xyz(ref_util, name):
    return ref_util.xyz2tensor(open(os.path.join(REF, "data", name)).read()).float()


def save(name, **arrays):
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **{k: (v.numpy() if isinstance(v, torch.Tensor) else np.asarray(v))
                                 for k, v in arrays.items()})
    print(f"wrote {path} ({os.path.getsize(path) / 1024:.1f} KiB)")


# ---------------------------------------------------------------- leaf functions
def gen_leaf(fu, util):
    out = {}
    sizes = [(1, 1), (1, 7), (7, 1), (33, 31), (255, 257), (256, 256), (1000, 333)]
    for k, (S, M) in enumerate(sizes):
        g = torch.Generator().manual_seed(1234 + k)
        src = rand_cloud(g, S)
        tgt = rand_cloud(g, M)
        out[f"c{k}_src"] = src
        out[f"c{k}_tgt"] = tgt
        out[f"c{k}_phi32"] = fu.potential(src, tgt)
        out[f"c{k}_E32"] = fu.field_grad(src, tgt)
        out[f"c{k}_E32_eps1e-6"] = fu.field_grad(src, tgt[:, :3], eps=1e-6)
        out[f"c{k}_phi64"] = fu.potential(src.double(), tgt.double())
        out[f"c{k}_E64"] = fu.field_grad(src.double(), tgt.double())
    out["n_cases"] = len(sizes)
    # recursion semantics (field_utils.py:23-44,73-94) exercised with a small max_pts
    g = torch.Generator().manual_seed(99)
    src = rand_cloud(g, 257)
    tgt = rand_cloud(g, 301)
    out["rec_src"], out["rec_tgt"] = src, tgt
    out["rec_phi32"] = fu.potential(src, tgt, max_pts=100)
    out["rec_E32"] = fu.field_grad(src, tgt, max_pts=100)
    out["rec_E64"] = fu.field_grad(src.double(), tgt.double(), max_pts=100)
    out["rec_phi64"] = fu.potential(src.double(), tgt.double(), max_pts=100)
    # coincident pairs: S == T (field_utils.py:98-106 masks them in field_grad;
    # potential has no mask and zeroes every target whose leaf total is NaN, :53-54)
    g = torch.Generator().manual_seed(7)
    cl = rand_cloud(g, 300)
    out["self_pc"] = cl
    out["self_E32"] = fu.field_grad(cl, cl)
    out["self_E64"] = fu.field_grad(cl.double(), cl.double())
    out["self_phi32"] = fu.potential(cl, cl)
    # coincident-pair exclusion equals removing the source: one
    # duplicate of target 5 appended to the sources
    src2 = torch.cat([rand_cloud(g, 64), cl[5:6]], dim=0)
    out["dup_src"] = src2
    out["dup_E32"] = fu.field_grad(src2, cl)
    out["dup_E64"] = fu.field_grad(src2.double(), cl.double())
    # analytic sphere: phi(centre) = N / r^2
    g = torch.Generator().manual_seed(11)
    sp = sphere_cloud(g, 8192, noise=0.0)
    probes = torch.tensor([[0.0, 0.0, 0.0], [0.1, -0.05, 0.02], [0.9, 0.9, 0.9]])
    out["sphere_pc"] = sp
    out["sphere_probes"] = probes
    out["sphere_phi32"] = fu.potential(sp, probes)
    out["sphere_phi64"] = fu.potential(sp.double(), probes.double())
    save("leaf", **out)


# ---------------------------------------------------------------- edge scores
def gen_edge(fu, util):
    out = {}
    g = torch.Generator().manual_seed(4321)
    pairs = [(120, 200), (100, 100), (257, 64), (1, 50)]
    for k, (s, t) in enumerate(pairs):
        S = sphere_cloud(g, s, r=0.3, noise=0.01, centre=(-0.1, 0.0, 0.0))
        T = sphere_cloud(g, t, r=0.3, noise=0.01, centre=(0.15, 0.05, 0.0))
        if k % 2:
            T[:, 3:] *= -1
        w, invw = fu.field_edge_calculator(S, T)
        out[f"e{k}_S"], out[f"e{k}_T"] = S, T
        out[f"e{k}_w"], out[f"e{k}_invw"] = np.asarray(w), np.asarray(invw)
        wb = fu.field_edge_calculator_bool(S, T)
        wc = fu.field_edge_calculator_count(S, T)
        out[f"e{k}_bool"] = np.asarray(wb)
        out[f"e{k}_count"] = np.asarray(wc)
        S64, T64 = S.double(), T.double()
        out[f"e{k}_w64"] = np.asarray(fu.field_edge_calculator(S64, T64)[0])
    out["n_cases"] = len(pairs)
    cl = sphere_cloud(g, 150, r=0.3, noise=0.01)
    out["selfall_pc"] = cl
    out["selfall_w"] = np.asarray(fu.self_interaction_all(cl))
    save("edge", **out)


# ---------------------------------------------------------------- orientation transfer
def gen_transfer(fu, util):
    out = {}
    g = torch.Generator().manual_seed(2024)
    ref = sphere_cloud(g, 500, r=0.4, noise=0.0)
    tgt = sphere_cloud(g, 700, r=0.4, noise=0.003)
    scr = (torch.rand(700, generator=g) < 0.5)
    tgt6 = tgt.clone()
    tgt6[scr, 3:] *= -1
    out["ref"], out["tgt3"], out["tgt6"] = ref, tgt[:, :3].clone(), tgt6.clone()
    out["out3"] = fu.reference_field(ref, tgt[:, :3].clone())
    out["out6"] = fu.reference_field(ref, tgt6.clone())
    save("transfer", **out)


# ---------------------------------------------------------------- lattice test
def gen_lattice(fu, util):
    out = {}
    out["grid"] = util.gen_grid()
    ok = load_xyz(util, "ok.xyz")
    ok_t, _ = util.Transform.trans(ok)
    out["ok_mean"] = fu.measure_mean_potential(ok_t)
    out["ok_mean_flipped"] = fu.measure_mean_potential(
        torch.cat([ok_t[:, :3], -ok_t[:, 3:]], dim=1))
    out["ok_phi"] = fu.potential(ok_t, util.gen_grid())
    out["ok_mean64"] = fu.measure_mean_potential(ok_t.double())
    g = torch.Generator().manual_seed(5)
    sp = sphere_cloud(g, 4000)
    out["sphere_pc"] = sp
    out["sphere_mean"] = fu.measure_mean_potential(sp)
    out["sphere_mean64"] = fu.measure_mean_potential(sp.double())
    save("lattice", **out)


# ---------------------------------------------------------------- drivers
class Recorder:
    """Wraps field_utils.field_grad to log (first source row) of each call."""

    def __init__(self, fu):
        self.fu = fu
        self.orig = fu.field_grad
        self.calls = []

    def __enter__(self):
        self.depth = 0

        def wrapped(sources, means, *a, **k):
            # field_grad recurses through the module attribute when a side exceeds max_pts: log top-level calls only
            if self.depth == 0:
                self.calls.append((sources[0].clone(), int(sources.shape[0]), int(means.shape[0])))
            self.depth += 1
            try:
                return self.orig(sources, means, *a, **k)
            finally:
                self.depth -= 1
        self.fu.field_grad = wrapped
        return self

    def __exit__(self, *exc):
        self.fu.field_grad = self.orig


def match_rows(xyz_all, rows):
    """index of each row (xyz) in xyz_all (exact match; clouds have no duplicates)."""
    lut = {tuple(p.tolist()): i for i, p in enumerate(xyz_all)}
    return np.array([lut[tuple(r[:3].tolist())] for r in rows], dtype=np.int64)


def scramble(pc, g, frac=0.5):
    flip = torch.rand(pc.shape[0], generator=g) < frac
    pc = pc.clone()
    pc[flip, 3:] *= -1
    return pc


def gen_points(fu, util):
    """cfg1: strongest_field_propagation_points on data/ok.xyz (field_utils.py:353-388)."""
    ok = load_xyz(util, "ok.xyz")
    ok_t, _ = util.Transform.trans(ok)
    g = torch.Generator().manual_seed(31337)
    pc_in = scramble(ok_t, g)
    out = {"pc_in": pc_in}
    for tag, n, diffuse in (("n512", 512, False), ("n2000", 2000, True), ("full", pc_in.shape[0], False)):
        pts = pc_in[:n].clone()
        t0 = time.time()
        with Recorder(fu) as rec:
            res = fu.strongest_field_propagation_points(pts, diffuse=diffuse, starting_point=0)
        dt = time.time() - t0
        order = match_rows(pc_in[:n, :3], [c[0] for c in rec.calls])
        # normal of the chosen point at the time it became a source (after the flip decision)
        src_n = torch.stack([c[0][3:] for c in rec.calls])
        flipped = ((src_n * pc_in[order, 3:]).sum(-1) < 0).numpy()
        out[f"{tag}_n"] = n
        out[f"{tag}_diffuse"] = diffuse
        out[f"{tag}_order"] = order.astype(np.int32)
        out[f"{tag}_flipped"] = flipped
        out[f"{tag}_normals"] = res[:, 3:]
        out[f"{tag}_mean_potential"] = fu.measure_mean_potential(res)
        out[f"{tag}_ref_seconds"] = dt
        print(f"points {tag}: N={n} ref CPU {dt:.1f}s")
    save("points_prop", **out)


def make_patches(util, pc, n_part, min_patch):
    idx, ijk = util._divide_pc(pc[:, :3], n_part, min_patch=min_patch)
    merged, _ = util.merge_nodes(pc, idx, ijk, min_patch)
    return merged


def patch_csr(patches):
    perm = torch.cat(patches).numpy().astype(np.int64)
    off = np.zeros(len(patches) + 1, dtype=np.int64)
    off[1:] = np.cumsum([len(p) for p in patches])
    return perm, off


def gen_patches(fu, util):
    """cfg2: strongest_field_propagation on data/fandisk.xyz (field_utils.py:286-348)."""
    fd = load_xyz(util, "fandisk.xyz")
    fd_t, _ = util.Transform.trans(fd)
    patches = make_patches(util, fd_t, 30, 100)
    P = len(patches)
    g = torch.Generator().manual_seed(777)
    # stand-in for the network stage: seeded per-patch flips, seeded weights
    pc_in = fd_t.clone()
    for p in patches:
        if torch.rand(1, generator=g).item() < 0.5:
            pc_in[p, 3:] *= -1
    weights = torch.rand(pc_in.shape[0], generator=g) * 0.9 + 0.05
    perm, off = patch_csr(patches)
    out = {"pc_in": pc_in, "weights": weights, "perm": perm, "offsets": off, "P": P}
    kept = [(i, p) for i, p in enumerate(patches) if i % 3 != 0]   # "network-kept" subset
    out["kept_ids"] = np.array([i for i, _ in kept], dtype=np.int64)
    first_pt = {int(p[0]): i for i, p in enumerate(patches)}
    for tag, diffuse, use_w in (("plain", False, False), ("diffuse_w", True, True)):
        pts = pc_in.clone()
        t0 = time.time()
        with Recorder(fu) as rec:
            fu.strongest_field_propagation(pts, kept, [p.clone() for p in patches], diffuse=diffuse,
                                           weights=weights.clone() if use_w else None)
        dt = time.time() - t0
        # sources of call k are pts[mask] for call 0 (sorted) and pts[patch] afterwards
        order = []
        for k, c in enumerate(rec.calls):
            row = match_rows(pc_in[:, :3], [c[0]])[0]
            if k == 0:
                pid = [i for i, p in enumerate(patches) if int(p.min()) == row][0]
            else:
                pid = first_pt[int(row)]
            order.append(pid)
        out[f"{tag}_order"] = np.array(order, dtype=np.int32)
        out[f"{tag}_normals"] = pts[:, 3:]
        out[f"{tag}_ref_seconds"] = dt
        out[f"{tag}_mean_potential"] = fu.measure_mean_potential(pts)
        print(f"patches {tag}: N={pts.shape[0]} P={P} ref CPU {dt:.1f}s")
    save("patch_prop", **out)


def gen_reps(fu, util):
    """cfg3 stand-in: strongest_field_propagation_reps (field_utils.py:207-282) on a noisy sphere pair."""
    g = torch.Generator().manual_seed(4242)
    a = sphere_cloud(g, 5000, r=0.33, noise=0.002, centre=(-0.12, 0.0, 0.0))
    b = sphere_cloud(g, 3000, r=0.2, noise=0.002, centre=(0.3, 0.1, 0.05))
    pc = torch.cat([a, b], dim=0)
    pc, _ = util.Transform.trans(pc)
    patches = make_patches(util, pc, 8, 60)
    P = len(patches)
    pc_in = pc.clone()
    for p in patches:
        if torch.rand(1, generator=g).item() < 0.5:
            pc_in[p, 3:] *= -1
    max_rep = 40
    reps = []
    for p in patches:
        pm = torch.randperm(p.shape[0], generator=g)
        reps.append((p[pm[:max_rep]], p[pm[max_rep:]]))
    rep_perm, rep_off = patch_csr([r for r, _ in reps])
    rest_perm, rest_off = patch_csr([r for _, r in reps])
    out = {"pc_in": pc_in, "rep_perm": rep_perm, "rep_off": rep_off,
           "rest_perm": rest_perm, "rest_off": rest_off, "P": P}
    first_pt = {int(r[0]): i for i, (r, _) in enumerate(reps)}
    for tag, diffuse in (("plain", False), ("diffuse", True)):
        pts = pc_in.clone()
        t0 = time.time()
        with Recorder(fu) as rec:
            fu.strongest_field_propagation_reps(pts, [(r.clone(), q.clone()) for r, q in reps], diffuse=diffuse)
        dt = time.time() - t0
        order = []
        for k, c in enumerate(rec.calls[:P]):       # last call is the reps -> non-reps field
            row = match_rows(pc_in[:, :3], [c[0]])[0]
            if k == 0:
                pid = [i for i, (r, _) in enumerate(reps) if int(r.min()) == row][0]
            else:
                pid = first_pt[int(row)]
            order.append(pid)
        out[f"{tag}_order"] = np.array(order, dtype=np.int32)
        out[f"{tag}_normals"] = pts[:, 3:]
        out[f"{tag}_ref_seconds"] = dt
        print(f"reps {tag}: N={pts.shape[0]} P={P} ref CPU {dt:.1f}s")
    save("reps_prop", **out)


def _sign_bits(normals, pc_in):
    """one bit per point: the output normal is the input normal negated"""
    return np.packbits(((normals * pc_in[:, 3:]).sum(-1) < 0).numpy())


def gen_large(fu, util):
    """cfg3 / cfg4 at the sizes BASELINE.md section 2 timed the reference (N = 60 000 / P = 48 representatives
    propagation, field_utils.py:207-282; 20 000 x 40 000 orientation transfer, :188-201).  The inputs are the
    seeded stand-ins of bench.py (regenerated by the tests; a checksum guards the generator); only decisions are
    stored: visit order, one flip bit per point, and the reference's wall times."""
    sys.path.insert(0, os.path.dirname(os.path.dirname(OUT)))
    import bench
    out = {}
    pc, reps = bench.cfg3_standin()
    out["cfg3_input_checksum"] = pc.double().sum().item()
    first_pt = {int(r[0]): i for i, (r, _) in enumerate(reps)}
    for tag, diffuse in (("plain", False), ("diffuse", True)):
        pts = pc.clone()
        t0 = time.time()
        with Recorder(fu) as rec:
            fu.strongest_field_propagation_reps(pts, [(r.clone(), q.clone()) for r, q in reps], diffuse=diffuse)
        dt = time.time() - t0
        order = []
        for k, c in enumerate(rec.calls[:len(reps)]):
            row = match_rows(pc[:, :3], [c[0]])[0]
            pid = [i for i, (r, _) in enumerate(reps) if int(r.min()) == row][0] if k == 0 else first_pt[int(row)]
            order.append(pid)
        out[f"cfg3_{tag}_order"] = np.array(order, dtype=np.int32)
        out[f"cfg3_{tag}_flip_bits"] = _sign_bits(pts[:, 3:], pc)
        out[f"cfg3_{tag}_ref_seconds"] = dt
        out[f"cfg3_{tag}_mean_potential"] = fu.measure_mean_potential(pts)
        print(f"large reps {tag}: N={pts.shape[0]} P={len(reps)} ref CPU {dt:.1f}s")
    ref, tgt = bench.cfg4_standin()
    out["cfg4_input_checksum"] = ref.double().sum().item() + tgt.double().sum().item()
    t0 = time.time()
    res6 = fu.reference_field(ref.clone(), tgt.clone())
    out["cfg4_ref_seconds"] = time.time() - t0
    out["cfg4_flip_bits"] = _sign_bits(res6[:, 3:], tgt)
    E = fu.field_grad(ref, tgt[:, :3])
    dots = (E * tgt[:, 3:]).sum(-1)
    out["cfg4_margin"] = (dots.abs() / (E.norm(dim=-1) + 1e-30)).numpy().astype(np.float32)   # |cos| of E vs n
    print(f"large transfer: {ref.shape[0]} x {tgt.shape[0]} ref CPU {out['cfg4_ref_seconds']:.1f}s")
    save("large", **out)


def gen_potsem(fu, util):
    """field_utils.potential on targets that coincide with sources (field_utils.py:53-54: the whole LEAF total of
    such a target becomes NaN -> 0, so the result depends on the max_pts recursion): small cases with several leaves,
    for the opt-in exact mode of the drop-in (field_utils.REFERENCE_NAN_SEMANTICS)."""
    g = torch.Generator().manual_seed(2718)
    out = {}
    cl = rand_cloud(g, 300)
    out["self_pc"] = cl
    for mp in (100, 64, 15000):
        out[f"self_phi_max{mp}"] = fu.potential(cl, cl, max_pts=mp)
    src, tgt = rand_cloud(g, 257), rand_cloud(g, 301)
    hit = torch.randperm(301, generator=g)[:40]
    tgt[hit, :3] = src[torch.randperm(257, generator=g)[:40], :3]
    tgt[hit[0], 0] = -0.0
    src[0, 0] = 0.0                                                    # +0 / -0 coordinates still coincide
    tgt[hit[0], 1:3] = src[0, 1:3]
    out["mix_src"], out["mix_tgt"], out["mix_hit"] = src, tgt, hit
    for mp in (100, 64, 15000):
        out[f"mix_phi_max{mp}"] = fu.potential(src, tgt, max_pts=mp)
    out["mix_phi_nonrecursive"] = fu.potential(src, tgt, recursive=False)
    save("potential_semantics", **out)


def gen_signatures(fu, util):
    """Names, parameter names and defaults of every hot-path function, so the
    drop-in's signatures can be compared with the reference's on a box without the reference."""
    import inspect
    import json
    names = ["measure_mean_potential", "potential", "field_grad", "field_edge_calculator",
             "field_edge_calculator_bool", "field_edge_calculator_count", "self_interaction",
             "self_interaction_all", "random_self_interaction", "reference_field",
             "strongest_field_propagation_reps", "strongest_field_propagation",
             "strongest_field_propagation_points"]
    out = {}
    for n in names:
        sig = inspect.signature(getattr(fu, n))
        out[n] = [[p.name, None if p.default is inspect.Parameter.empty else repr(p.default)]
                  for p in sig.parameters.values()]
    path = os.path.join(OUT, "signatures.json")
    json.dump({"field_utils": out, "reexports": ["torch", "np", "util"]}, open(path, "w"), indent=1)
    print(f"wrote {path}")


def gen_xie(fu, util):
    """the "xie" reflected-normal interaction (field_utils.py:431-469,509-519) and
    the in-order propagation over T visiting orders (:569-605).  float32 = what the reference computes,
    float64 = the same functions on .double() inputs (accuracy referee)."""
    g = torch.Generator().manual_seed(4242)
    out = {}
    src, tgt = scramble(sphere_cloud(g, 200), g), scramble(sphere_cloud(g, 120), g)
    tgt[:7] = src[:7]                                          # coincident pairs: the un-normalised branch (:456-460)
    out["field_C3"] = fu.xie_field(src, tgt, 1e-2, C=3)
    for C in (3, 2):
        out[f"inter_C{C}"] = fu.xie_intersaction(src, tgt, 1e-2, -1, C)
        out[f"inter64_C{C}"] = fu.xie_intersaction(src.double(), tgt.double(), 1e-2, -1, C)
    out["src"], out["tgt"] = src, tgt
    # knn mask (:447-455): each source keeps its k nearest targets
    out["inter_knn16"] = fu.xie_intersaction(src, tgt, 1e-2, 16, 3)
    # in-order propagation on the shipped cloud: T visiting orders (seeded permutations; the reference builds BFS
    # routes of a kNN graph, graph.py:380-392 -- any order exercises the same arithmetic)
    pc = scramble(load_xyz(util, "ok.xyz")[:1500], g)
    N = pc.shape[0]
    orders = np.stack([torch.randperm(N, generator=g).numpy() for _ in range(5)])
    out["prop_pc"], out["prop_orders"] = pc, orders
    for diffuse in (False, True):
        for dt, tag in ((torch.float32, "f32"), (torch.float64, "f64")):
            flips = fu.xie_propagation_points_in_order(pc.clone().to(dt), 1e-2, orders, diffuse=diffuse, knn_mask=-1, C=3)
            out[f"prop_flips_{tag}_diffuse{int(diffuse)}"] = flips
    import inspect
    import json
    sig = {n: [[p.name, None if p.default is inspect.Parameter.empty else repr(p.default)]
               for p in inspect.signature(getattr(fu, n)).parameters.values()]
           for n in ("xie_field", "xie_intersaction", "xie_propagation_points_in_order")}
    json.dump(sig, open(os.path.join(OUT, "xie_signatures.json"), "w"), indent=1)
    save("xie", **out)


def gen_partition(fu, util):
    """voxel partition util._divide_pc (util.py:110-150) and the small-patch merge
    util.merge_nodes (util.py:448-492), on the transformed demo clouds with the demos' own flags."""
    import inspect
    import json
    out = {}
    cases = [("fandisk", "fandisk.xyz", 30, 100), ("ok", "ok.xyz", 41, 100), ("hand", "hand.xyz", 10, 40)]
    for tag, name, n_part, min_patch in cases:
        pc, _ = util.Transform.trans(load_xyz(util, name))
        if tag == "hand":
            pc[::97, 0] = 7.0                                      # outside the (-1.5, 1.5] box: dropped by the reference
            pc[5, :3] = torch.tensor([-1.5 + 3.0 / n_part, 0.0, 0.0])   # exactly on a cell boundary (belongs to the lower cell)
        t0 = time.time()
        idx, ijk = util._divide_pc(pc[:, :3], n_part, min_patch=min_patch)
        dt = time.time() - t0
        out[f"{tag}_pc"] = pc
        out[f"{tag}_n_part"], out[f"{tag}_min_patch"] = n_part, min_patch
        out[f"{tag}_cell_perm"], out[f"{tag}_cell_offsets"] = patch_csr([c[0] for c in idx])
        out[f"{tag}_cell_ijk"] = torch.stack([c[0] for c in ijk]).numpy()
        out[f"{tag}_divide_seconds"] = dt
        t0 = time.time()
        merged, merged_ijk = util.merge_nodes(pc, idx, ijk, min_patch)
        out[f"{tag}_merge_seconds"] = time.time() - t0
        out[f"{tag}_merged_perm"], out[f"{tag}_merged_offsets"] = patch_csr(merged)
        out[f"{tag}_merged_ijk"] = np.concatenate([torch.stack(c).numpy() for c in merged_ijk])
        out[f"{tag}_merged_ijk_offsets"] = np.concatenate([[0], np.cumsum([len(c) for c in merged_ijk])])
        print(f"partition {tag}: N={pc.shape[0]} n_part={n_part}: {len(idx)} cells in {dt:.1f}s, {len(merged)} patches")
    sig = {n: [[p.name, None if p.default is inspect.Parameter.empty else repr(p.default)]
               for p in inspect.signature(getattr(util, n)).parameters.values()]
           for n in ("_divide_pc", "divide_pc", "merge_nodes")}
    json.dump(sig, open(os.path.join(OUT, "partition_signatures.json"), "w"), indent=1)
    save("partition", **out)


def gen_graph(fu, util):
    """the objective the reference hands to Gurobi (graph.cal_loss, graph.py:431-442),
    the score container (BiEdge / BidGraph.to_matrix, graph.py:121-152) and the BFS routes of LinkedListGraph
    (graph.py:257-318).  Gurobi itself is absent, so the maximiser is not recorded here (oracle: enumeration)."""
    import inspect
    import json
    cwd = os.getcwd()
    os.chdir(REF)
    try:
        import graph as ref_graph
    finally:
        os.chdir(cwd)
    rng = np.random.default_rng(20260)
    out = {}
    for tag, n, sym in (("sym5", 5, True), ("sym12", 12, True), ("gen9", 9, False), ("sym40", 40, True)):
        if sym:
            W = np.triu(rng.normal(size=(n, n)), 1) * (rng.random((n, n)) < 0.5)
            A = W + W.T
            B = -A
        else:
            A = rng.normal(size=(n, n))
            B = rng.normal(size=(n, n))
        X = (rng.random((16, n)) < 0.5).astype(np.float64)
        out[f"{tag}_A"], out[f"{tag}_B"], out[f"{tag}_X"] = A, B, X
        out[f"{tag}_loss"] = np.array([ref_graph.cal_loss(x, A, B) for x in X])
    # score container
    G = ref_graph.BidGraph()
    G.V = list(range(7))
    edges = [(0, 1, 0.5, -0.5), (1, 2, -1.25, 1.25), (2, 6, 3.0, -3.0), (3, 4, 0.125, -0.125), (0, 5, -2.0, 2.0)]
    for u, v, w, invw in edges:
        G.E.append(ref_graph.BiEdge(u, v, np.float32(w), np.float32(invw)))
    A, B = G.to_matrix()
    out["container_edges"] = np.array(edges)
    out["container_A"], out["container_B"] = A, B
    # BFS routes
    n = 60
    L = ref_graph.LinkedListGraph(n)
    pairs = []
    for i in range(n):
        for j in rng.choice(n, size=3, replace=False):
            if i != int(j) and i // 20 == int(j) // 20:       # three components
                w = float(rng.random())
                L.add_edge(i, int(j), w)
                L.add_edge(int(j), i, w)
                pairs += [(i, int(j), w), (int(j), i, w)]
    out["bfs_n"] = n
    out["bfs_edges"] = np.array(pairs)
    out["bfs_route_0"] = np.array(L.get_bfs_route(0))
    out["bfs_route_33"] = np.array(L.get_bfs_route(33))
    out["bfs_weighted_route_0"] = np.array(L.get_weighted_bfs_route(0))
    out["bfs_weighted_route_45"] = np.array(L.get_weighted_bfs_route(45))
    # BFS-tree propagation + flip vote (field_utils.py:654-709).  Its MIQP (field_utils.py:618-643, objective
    # -cal_loss) needs Gurobi: replaced HERE ONLY by enumeration of the same objective (first minimiser in
    # ascending integer order), so everything around the solve -- seeds, kNN graph, BFS routes, chains, counts,
    # majority vote -- is the reference's own.  x and 1 - x tie, so the result is defined up to a global sign.
    votes = {}

    def enum_miqp(A, B):
        n = len(A)
        best, arg = None, None
        for s_ in range(1 << n):
            x = np.array([(s_ >> i) & 1 for i in range(n)], dtype=np.float64)
            v = fu.cal_loss(x, A, B)
            if best is None or v < best:
                best, arg = v, x
        votes["A"], votes["B"], votes["x"], votes["min_loss"] = np.array(A), np.array(B), arg, best
        return arg
    fu.MIQP = enum_miqp
    g = torch.Generator().manual_seed(977)
    pc = scramble(load_xyz(util, "ok.xyz")[:1500], g)
    pc, _ = util.Transform.trans(pc)
    for times in (1, 5):
        work = pc.clone()
        fu.xie_propagation_points_onbfstree(work, 1e-2, diffuse=True, starting_point=0, times=times, k=10, treshold=0.1)
        out[f"tree_result_t{times}"] = work
        out[f"tree_vote_A_t{times}"], out[f"tree_vote_B_t{times}"] = votes["A"], votes["B"]
        out[f"tree_vote_x_t{times}"], out[f"tree_vote_min_t{times}"] = votes["x"], votes["min_loss"]
    out["tree_pc"] = pc
    sig = {}
    fsig = {n: [[p.name, None if p.default is inspect.Parameter.empty else repr(p.default)]
                for p in inspect.signature(getattr(fu, n)).parameters.values()]
            for n in ("cal_loss", "xie_propagation_points_onbfstree")}
    fsig["MIQP"] = [["A", None], ["B", None]]                   # field_utils.py:618 (replaced above, so written out)
    json.dump(fsig, open(os.path.join(OUT, "graph_field_utils_signatures.json"), "w"), indent=1)
    for name in ("cal_loss", "MIQP"):
        sig[name] = [[p.name, None if p.default is inspect.Parameter.empty else repr(p.default)]
                     for p in inspect.signature(getattr(ref_graph, name)).parameters.values()]
    for cls, meths in (("BiEdge", ["__init__"]), ("BidGraph", ["__init__", "to_matrix", "add_edge"]),
                       ("LinkedListGraph", ["__init__", "add_edge", "get_bfs_route", "get_weighted_bfs_route"])):
        for m in meths:
            sig[f"{cls}.{m}"] = [[p.name, None if p.default is inspect.Parameter.empty else repr(p.default)]
                                 for p in inspect.signature(getattr(getattr(ref_graph, cls), m)).parameters.values()]
    json.dump(sig, open(os.path.join(OUT, "graph_signatures.json"), "w"), indent=1)
    save("graph", **out)


def gen_patchgraph(fu, util):
    """util.divide_pc_to_graph (util.py:160-212) with its grid partition
    (_lzd_divide_pc, :85-108), merge (lzd_merge_nodes, :418-447), per-patch orientation and flip scores.
    Shim outside the math: util.py:162 calls `util.timer_factory()` without importing itself (NameError at
    HEAD) -> `util.util = util`.  The per-patch estimator is graph_dipole.st_propagation_points_file
    (graph_dipole.py:27-36) restated here from the reference's own functions, because importing
    graph_dipole.py creates directories (graph_dipole.py:9-25)."""
    util.util = util

    def st_points(pc, verbose=False):
        pc, tr = util.Transform.trans(pc)
        fu.strongest_field_propagation_points(pc, diffuse=True, starting_point=0, verbose=verbose)
        if fu.measure_mean_potential(pc) < 0:
            pc[:, 3:] *= -1
        return tr.inverse(pc)
    g = torch.Generator().manual_seed(31337)
    out = {}
    for tag, name, n_part, min_patch, estimator in (("hand", "hand.xyz", 10, 0, st_points),
                                                    ("fandisk", "fandisk.xyz", 8, 60, None)):
        pc, _ = util.Transform.trans(scramble(load_xyz(util, name), g))
        out[f"{tag}_pc"] = pc.clone()
        out[f"{tag}_n_part"], out[f"{tag}_min_patch"] = n_part, min_patch
        idx, ijk = util._lzd_divide_pc(pc, n_part, min_patch=min_patch)
        out[f"{tag}_cell_perm"], out[f"{tag}_cell_offsets"] = patch_csr([c[0] for c in idx])
        out[f"{tag}_cell_ijk"] = torch.stack([c[0] for c in ijk]).numpy()
        m_idx, m_ijk, m_src = util.lzd_merge_nodes(pc, idx, ijk, min_patch)
        out[f"{tag}_merged_perm"], out[f"{tag}_merged_offsets"] = patch_csr(m_idx)
        out[f"{tag}_merged_ijk"] = torch.stack(m_ijk).numpy()
        out[f"{tag}_source_ijk"] = np.concatenate([torch.stack(c).numpy() for c in m_src])
        out[f"{tag}_source_offsets"] = np.concatenate([[0], np.cumsum([len(c) for c in m_src])])
        work = pc.clone()
        t0 = time.time()
        G, index = util.divide_pc_to_graph(work, n_part=n_part, min_patch=min_patch,
                                           edge_calculator=fu.field_edge_calculator, point_estimator=estimator)
        out[f"{tag}_graph_seconds"] = time.time() - t0
        out[f"{tag}_graph_pc"] = work
        out[f"{tag}_graph_perm"], out[f"{tag}_graph_offsets"] = patch_csr(index)
        out[f"{tag}_graph_edges"] = np.array([[e.u, e.v] for e in G.E])
        out[f"{tag}_graph_w"] = np.array([float(e.w) for e in G.E], dtype=np.float32)
        out[f"{tag}_graph_invw"] = np.array([float(e.invw) for e in G.E], dtype=np.float32)
        A, B = G.to_matrix()
        out[f"{tag}_A"], out[f"{tag}_B"] = A, B
        print(f"patchgraph {tag}: {len(index)} patches, {len(G.E)} edges, {out[f'{tag}_graph_seconds']:.1f}s")
    import inspect
    import json
    sig = {n: [[p.name, None if p.default is inspect.Parameter.empty else repr(p.default)]
               for p in inspect.signature(getattr(util, n)).parameters.values()]
           for n in ("_lzd_divide_pc", "lzd_merge_nodes", "divide_pc_to_graph", "npxyz2tensor", "xyz2tensor", "export_pc",
                     "estimate_normals", "estimate_normals_torch", "orient_center", "timer_factory")}
    json.dump(sig, open(os.path.join(OUT, "util_signatures.json"), "w"), indent=1)
    save("patchgraph", **out)


def gen_textio(fu, util):
    """util.xyz2tensor (util.py:53-69) and util.export_pc (util.py:46-51) on a text
    with the cases the format allows: 3- and 6-column lines, blank / short / long lines, tabs and CR at the line
    ends, 'nan' lines, exponents, 17-digit reprs, signed zeros, long literals, inf."""
    import tempfile
    rng = np.random.default_rng(4711)
    lines = ["0.1 0.2 0.3 1 0 0", "  -1.5e-3 2.5E+2 .5 0. -0.0 +7  ", "1 2 3", "", "1 2", "1 2 3 4", "1 2 3 4 5 6 7",
             "0.1 nan 0.3", "3 4 nan 0 0 1", "NaN 1 2", "1\t2\t3", "\t9 8 7\r", "0.30000001192092896 1e-5 1e22",
             "123456789012345678901234567890 0.000000000000000000000123 1e-40", "inf -inf Infinity",
             "4.9406564584124654e-40 3.4028234663852886e+38 1.1754943508222875e-38",
             "16777217 16777219 0.100000001490116119384765625", "1e0 1E1 1e+2", "5e-46 7.006492321624085e-46 1e39",
             "9007199254740993 9007199254740992.5 0.5000000298023224"]
    vals = np.concatenate([rng.normal(size=300), rng.normal(size=300) * 1e-6, rng.normal(size=300) * 1e7,
                           rng.integers(-50, 50, 300).astype(np.float64)]).astype(np.float32)
    rng.shuffle(vals)
    for i in range(0, 1200, 6):
        lines.append(' '.join(repr(float(v)) for v in vals[i:i + 6]))
    for i in range(0, 600, 3):
        lines.append(' '.join(f"{float(v):.9g}" for v in vals[i:i + 3]))
    txt = "\n".join(lines) + "\n"
    out = {"text": np.frombuffer(txt.encode(), dtype=np.uint8)}
    out["parsed_append"] = util.xyz2tensor(txt, append_normals=True)
    six_only = "\n".join(l for l in lines if len(l.strip().split(' ')) != 3)
    out["text_six"] = np.frombuffer(six_only.encode(), dtype=np.uint8)
    out["parsed_six_noappend"] = util.xyz2tensor(six_only, append_normals=False)
    three_only = "\n".join(l for l in lines if len(l.strip().split(' ')) != 6)
    out["text_three"] = np.frombuffer(three_only.encode(), dtype=np.uint8)
    out["parsed_three_noappend"] = util.xyz2tensor(three_only, append_normals=False)
    # export: float32 values of every magnitude, special values, powers of two and ten
    bits = rng.integers(0, 2 ** 32, 6 * 4000, dtype=np.uint64).astype(np.uint32)
    ex = bits.view(np.float32).copy()
    ex = ex[~np.isnan(ex)][:6 * 3000]
    special = np.array([0.0, -0.0, 1.0, -1.0, 0.1, 0.5, 1e-5, 1e-4, 1e16, 1e15, 9.999999e15, 1.0000001e16, 123456.0,
                        np.inf, -np.inf, 3.4028235e38, 1e-45, 1.1754944e-38, 2.0 ** -126, 2.0 ** 24, 2.0 ** 60,
                        1e-3, 0.3, 16777216.0, 1e10, 1e22, 1e23, 5e-324 + 1e-40, 0.001, 1234.5], dtype=np.float32)
    ex = np.concatenate([special, ex]).astype(np.float32)
    ex = ex[:(len(ex) // 6) * 6].reshape(-1, 6)
    pc = torch.from_numpy(ex)
    with tempfile.TemporaryDirectory() as d:
        dest = os.path.join(d, "out.xyz")
        util.export_pc(pc.transpose(0, 1), dest)
        written = open(dest, "rb").read()
    out["export_pc"] = ex
    out["export_text"] = np.frombuffer(written, dtype=np.uint8)
    # the demo cloud round trip
    ok_txt = open(os.path.join(REF, "data", "ok.xyz")).read()
    ok = util.xyz2tensor(ok_txt)
    out["ok_rows"] = ok[:2000]
    out["ok_n"] = ok.shape[0]
    save("textio", **out)


def gen_parser(fu, util):
    """dipole_api.get_parser (dipole_api.py:93-132) and options.get_parser (options.py:5-19): flags, types and
    defaults.  dipole_api.py cannot be imported here (inference_utils -> torch_geometric), so its parser section
    (module lines 89-132: imports, path constants, get_parser) is executed on its own."""
    import json
    src = open(os.path.join(REF, "dipole_api.py")).read().split("\n")
    start = next(i for i, l in enumerate(src) if l.startswith("import argparse"))
    end = next(i for i, l in enumerate(src) if l.startswith("if __name__"))
    ns = {}
    exec("\n".join(src[start:end]), ns)

    def describe(parser):
        rows = []
        for a in parser._actions:
            if a.dest == "help":
                continue
            rows.append({"flags": a.option_strings, "dest": a.dest, "type": getattr(a.type, "__name__", None),
                         "default": repr(a.default), "nargs": a.nargs, "action": type(a).__name__,
                         "required": a.required})
        return rows
    sys.path.insert(0, REF)
    import options as ref_options
    out = {"dipole_api.get_parser": describe(ns["get_parser"]()),
           "options.get_parser": describe(ref_options.get_parser())}
    json.dump(out, open(os.path.join(OUT, "parser_signatures.json"), "w"), indent=1)
    print("wrote parser_signatures.json")


GENERATORS = {"signatures": gen_signatures, "leaf": gen_leaf, "edge": gen_edge, "transfer": gen_transfer, "lattice": gen_lattice,
              "points": gen_points, "patches": gen_patches, "reps": gen_reps, "xie": gen_xie, "partition": gen_partition,
              "graph": gen_graph, "patchgraph": gen_patchgraph, "textio": gen_textio, "parser": gen_parser,
              "large": gen_large, "potsem": gen_potsem}

if __name__ == "__main__":
    torch.manual_seed(0)
    np.random.seed(0)
    fu, util = import_reference()
    which = sys.argv[1:] or list(GENERATORS)
    for name in which:
        GENERATORS[name](fu, util)
