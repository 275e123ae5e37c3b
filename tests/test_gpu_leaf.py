"""GPU parity of the leaf functions through the reference-facing
Python mirror (field_utils.py -> ctypes -> libdipole.so C ABI).

Stated FP32 tolerance, per target, against the float64 oracle / the reference's
float64 golden outputs:
    |E_gpu - E_ref|     <= 1e-4 |E_ref|   + 1e-6 max_m |E_ref|
    |phi_gpu - phi_ref| <= 1e-4 |phi_ref| + 1e-5 max_m |phi_ref|
and the CUDA path may be no worse than 4x the reference's own float32-vs-float64 error.
Signs and decisions (orientation transfer, score signs, lattice sign) must be identical.
"""
import os
import threading

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

RTOL, ATOL_E, ATOL_PHI = 1e-4, 1e-6, 1e-5


def check_E(E, ref, ctx=""):
    E = np.asarray(E, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert E.shape == ref.shape, ctx
    err = np.linalg.norm(E - ref, axis=-1)
    mag = np.linalg.norm(ref, axis=-1)
    bound = RTOL * mag + ATOL_E * (mag.max() if mag.size else 0.0)
    assert np.all(err <= bound), f"{ctx}: worst excess {np.max(err - bound):.3e} (max |E| {mag.max():.3e})"


def check_phi(phi, ref, ctx=""):
    phi = np.asarray(phi, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert phi.shape == ref.shape, ctx
    bound = RTOL * np.abs(ref) + ATOL_PHI * (np.abs(ref).max() if ref.size else 0.0)
    assert np.all(np.abs(phi - ref) <= bound), f"{ctx}: worst excess {np.max(np.abs(phi - ref) - bound):.3e}"


def cloud(n, seed, cols=6):
    g = torch.Generator().manual_seed(seed)
    xyz = torch.rand(n, 3, generator=g) - 0.5
    if cols == 3:
        return xyz.float()
    nr = torch.randn(n, 3, generator=g)
    nr = nr / nr.norm(dim=-1, keepdim=True)
    return torch.cat([xyz, nr], 1).float()


@pytest.fixture(scope="module")
def fu():
    import field_utils
    return field_utils


@pytest.fixture()
def force(monkeypatch):
    """Force the kernel variant (targets per thread / source splits) through the tuning overrides."""
    def _set(tp=None, nsplit=None):
        for k, v in (("DNP_FORCE_TP", tp), ("DNP_FORCE_NSPLIT", nsplit)):
            if v is None:
                monkeypatch.delenv(k, raising=False)
            else:
                monkeypatch.setenv(k, str(v))
    return _set


# ------------------------------------------------------------------ golden fixtures of the reference
def test_golden_leaf_cases_cuda_tensors(fu, golden, dev):
    g = golden("leaf")
    for k in range(int(g["n_cases"])):
        src = torch.from_numpy(g[f"c{k}_src"]).to(dev)
        tgt = torch.from_numpy(g[f"c{k}_tgt"]).to(dev)
        phi = fu.potential(src, tgt)
        E = fu.field_grad(src, tgt)
        E6 = fu.field_grad(src, tgt[:, :3], eps=1e-6)            # strided 3-column view, point-propagation eps
        assert phi.device == src.device and phi.dtype == torch.float32 and phi.shape == (tgt.shape[0],)
        assert E.shape == (tgt.shape[0], 3)
        check_phi(phi.cpu().numpy(), g[f"c{k}_phi64"], f"case {k}")
        check_E(E.cpu().numpy(), g[f"c{k}_E64"], f"case {k}")
        check_E(E6.cpu().numpy(), g[f"c{k}_E32_eps1e-6"], f"case {k} eps=1e-6")
        # no worse than 4x the reference's own float32 error
        ref_err = np.abs(g[f"c{k}_E32"] - g[f"c{k}_E64"]).max()
        our_err = np.abs(E.cpu().numpy() - g[f"c{k}_E64"]).max()
        assert our_err <= 4 * ref_err + 1e-6 * np.abs(g[f"c{k}_E64"]).max(), k


def test_golden_leaf_cases_host_tensors(fu, golden):
    """CPU tensors in, CPU tensors out: the host-buffer ABI entry points (copies inside the call)."""
    g = golden("leaf")
    for k in range(int(g["n_cases"])):
        src, tgt = torch.from_numpy(g[f"c{k}_src"]), torch.from_numpy(g[f"c{k}_tgt"])
        phi, E = fu.potential(src, tgt), fu.field_grad(src, tgt)
        assert not phi.is_cuda and not E.is_cuda
        check_phi(phi.numpy(), g[f"c{k}_phi64"], f"case {k}")
        check_E(E.numpy(), g[f"c{k}_E64"], f"case {k}")


def test_golden_recursive_case(fu, golden, dev):
    g = golden("leaf")
    src, tgt = torch.from_numpy(g["rec_src"]).to(dev), torch.from_numpy(g["rec_tgt"]).to(dev)
    check_E(fu.field_grad(src, tgt, max_pts=100).cpu().numpy(), g["rec_E64"])
    check_phi(fu.potential(src, tgt, max_pts=100).cpu().numpy(), g["rec_phi64"])


def test_coincident_pairs(fu, golden, orc, dev):
    g = golden("leaf")
    cl = torch.from_numpy(g["self_pc"]).to(dev)
    E = fu.field_grad(cl, cl).cpu().numpy()
    assert np.isfinite(E).all()
    check_E(E, g["self_E64"], "S == T")
    dup = torch.from_numpy(g["dup_src"]).to(dev)
    check_E(fu.field_grad(dup, cl).cpu().numpy(), g["dup_E64"], "duplicated source")
    # potential: documented divergence -- only the coincident pair is dropped (oracle mode 1)
    phi = fu.potential(cl, cl).cpu().numpy()
    assert np.isfinite(phi).all()
    check_phi(phi, orc.potential(g["self_pc"], g["self_pc"], drop_coincident=True), "S == T")


def test_analytic_sphere(fu, golden, dev):
    g = golden("leaf")
    sp = torch.from_numpy(g["sphere_pc"]).to(dev)
    probes = torch.from_numpy(g["sphere_probes"]).to(dev)
    phi = fu.potential(sp, probes).cpu().numpy()
    check_phi(phi, g["sphere_phi64"])
    r = np.linalg.norm(g["sphere_pc"][:, :3], axis=1).mean()
    assert abs(phi[0] - sp.shape[0] / r ** 2) / phi[0] < 1e-4


# ------------------------------------------------------------------ seeded sizes vs the oracle, every kernel variant
SIZES = [(1, 1), (7, 1), (1, 7), (255, 257), (256, 256), (257, 255), (1000, 3333), (15000, 15001), (15001, 2048),
         (65536, 1024), (1024, 65536)]


@pytest.mark.parametrize("S,M", SIZES)
def test_sizes_vs_oracle(fu, orc, dev, S, M):
    src, tgt = cloud(S, 1000 + S), cloud(M, 2000 + M)
    E = fu.field_grad(src.to(dev), tgt.to(dev)).cpu().numpy()
    phi = fu.potential(src.to(dev), tgt.to(dev)).cpu().numpy()
    check_E(E, orc.field_grad(src.numpy(), tgt.numpy()), f"{S}x{M}")
    check_phi(phi, orc.potential(src.numpy(), tgt.numpy()), f"{S}x{M}")


@pytest.mark.parametrize("tp", [1, 2, 4])
@pytest.mark.parametrize("nsplit", [1, 3, 7])
def test_kernel_variants(fu, orc, dev, force, tp, nsplit):
    src, tgt = cloud(5000, 5), cloud(3001, 6)
    force(tp, nsplit)
    E = fu.field_grad(src.to(dev), tgt.to(dev)).cpu().numpy()
    phi = fu.potential(src.to(dev), tgt.to(dev)).cpu().numpy()
    check_E(E, orc.field_grad(src.numpy(), tgt.numpy()), f"tp={tp} nsplit={nsplit}")
    check_phi(phi, orc.potential(src.numpy(), tgt.numpy()), f"tp={tp} nsplit={nsplit}")


@pytest.mark.parametrize("nsplit", [None, 1, 3, 7])
@pytest.mark.parametrize("S,M", [(5000, 3001), (150, 11031), (63, 64), (64, 1), (65, 700)])
def test_small_source_tiles(fu, orc, dev, force, monkeypatch, S, M, nsplit):
    """The 64-source-tile variant small calls take (a patch against a small cloud), forced at every size."""
    monkeypatch.setenv("DNP_FORCE_TILE", "64")
    force(None, nsplit)
    src, tgt = cloud(S, 21), cloud(M, 22)
    E = fu.field_grad(src.to(dev), tgt.to(dev)).cpu().numpy()
    phi = fu.potential(src.to(dev), tgt.to(dev)).cpu().numpy()
    check_E(E, orc.field_grad(src.numpy(), tgt.numpy()), f"tile 64 nsplit={nsplit}")
    check_phi(phi, orc.potential(src.numpy(), tgt.numpy()), f"tile 64 nsplit={nsplit}")


@pytest.mark.parametrize("S,M", [(1000, 250000), (2000, 65536), (500, 300000), (150, 11031)])
def test_planned_splits_match_the_single_pass(fu, dev, force, S, M):
    """Mid-size calls (a patch against every representative): the plan's (targets per thread, splits) choice is a
    scheduling decision only -- same field as the unsplit TP4 kernel up to float32 summation order."""
    src, tgt = cloud(S, 11).to(dev), cloud(M, 12).to(dev)
    force(None, None)
    E = fu.field_grad(src, tgt)
    force(4, 1)
    E_ref = fu.field_grad(src, tgt)
    scale = E_ref.abs().max().item()
    assert (E - E_ref).abs().max().item() <= 2e-5 * scale


def test_workspace_sized_for_another_source_count(fu, dev, monkeypatch):
    """The greedy drivers size one workspace for their largest patch: a call whose plan wants more splits than the
    workspace has room for keeps the splits that fit instead of failing."""
    import dnp_native as dn
    S, M = 2000, 65536                                            # plan: 4 splits
    src, tgt = cloud(S, 13).to(dev), cloud(M, 14).to(dev)
    E_ref = fu.field_grad(src, tgt)
    packed = 2048 * 48                                            # 3 float4 per packed source
    for room_splits in (0, 2.5):
        nbytes = packed + int(room_splits * M * 12)
        monkeypatch.setattr(dn, "_workspace",
                            lambda s, m, d, nb=nbytes: (torch.empty(nb, dtype=torch.uint8, device=d), nb))
        E = fu.field_grad(src, tgt)
        assert (E - E_ref).abs().max().item() <= 2e-5 * E_ref.abs().max().item()
    monkeypatch.setattr(dn, "_workspace", lambda s, m, d: (torch.empty(1024, dtype=torch.uint8, device=d), 1024))
    with pytest.raises(Exception, match="workspace too small"):
        fu.field_grad(src, tgt)


def test_surface_cloud_with_near_pairs(fu, orc, dev):
    """Noisy-sphere cloud evaluated on itself shifted by 1e-4: near-singular pairs, eps matters."""
    g = torch.Generator().manual_seed(77)
    d = torch.randn(6000, 3, generator=g)
    d = d / d.norm(dim=-1, keepdim=True)
    src = torch.cat([d * 0.45 + torch.randn(6000, 3, generator=g) * 0.002, d], 1).float()
    tgt = src[:, :3] + 1e-4
    for eps in (1e-5, 1e-6, 0.0):
        E = fu.field_grad(src.to(dev), tgt.to(dev), eps=eps).cpu().numpy()
        check_E(E, orc.field_grad(src.numpy(), tgt.numpy(), eps=eps), f"eps={eps}")


def test_views_and_strides(fu, orc, dev):
    big = torch.rand(700, 9, generator=torch.Generator().manual_seed(3)).to(dev) - 0.5
    src = big[:, 2:8]                 # column-offset view, row stride 9
    tgt = big[::2, 1:4]               # every second row
    E = fu.field_grad(src, tgt).cpu().numpy()
    check_E(E, orc.field_grad(src.cpu().numpy(), tgt.cpu().numpy()))
    phi = fu.potential(src, tgt).cpu().numpy()
    check_phi(phi, orc.potential(src.cpu().numpy(), tgt.cpu().numpy()))


def test_empty_inputs(fu, dev):
    src, tgt = cloud(10, 1).to(dev), cloud(5, 2).to(dev)
    assert fu.field_grad(src, tgt[:0]).shape == (0, 3)
    assert fu.potential(src, tgt[:0]).shape == (0,)
    assert torch.equal(fu.field_grad(src[:0], tgt), torch.zeros(5, 3, device=dev))      # sum over no sources
    assert torch.equal(fu.potential(src[:0], tgt), torch.zeros(5, device=dev))


def test_nonfinite_inputs_are_zeroed(fu, dev):
    """field_utils.py:114-115 / :53-54: NaN/Inf totals come back as 0."""
    src, tgt = cloud(64, 1).to(dev), cloud(32, 2).to(dev)
    src[3, 0] = float("nan")
    E, phi = fu.field_grad(src, tgt), fu.potential(src, tgt)
    assert torch.isfinite(E).all() and torch.isfinite(phi).all()
    assert (E == 0).all() and (phi == 0).all()


def test_errors(fu, dev):
    import dnp_native
    a = cloud(8, 1).to(dev)
    with pytest.raises(TypeError):
        fu.field_grad(a.double(), a)
    with pytest.raises(ValueError):
        fu.field_grad(a, a.cpu())
    with pytest.raises(ValueError):
        fu.field_grad(a[:, :5], a)
    with pytest.raises(dnp_native.DipoleError):
        fu.field_grad(a, a, eps=-1.0)


def test_deterministic(fu, dev):
    src, tgt = cloud(20000, 1).to(dev), cloud(9000, 2).to(dev)
    a, b = fu.field_grad(src, tgt), fu.field_grad(src, tgt)
    assert torch.equal(a, b)                                    # fixed reduction order, no float atomics


def test_concurrent_host_threads(fu, orc, dev):
    """The reference calls this path from many Python threads (util.py:187-196,308-327)."""
    jobs = [(cloud(700 + 13 * i, i), cloud(500 + 7 * i, 100 + i)) for i in range(8)]
    want = [orc.field_grad(s.numpy(), t.numpy()) for s, t in jobs]
    got = [None] * len(jobs)

    def run(i):
        stream = torch.cuda.Stream(device=dev)
        with torch.cuda.stream(stream):
            for _ in range(5):
                got[i] = fu.field_grad(jobs[i][0].to(dev), jobs[i][1].to(dev))
            stream.synchronize()

    threads = [threading.Thread(target=run, args=(i,)) for i in range(len(jobs))]
    [t.start() for t in threads]
    [t.join() for t in threads]
    for i in range(len(jobs)):
        check_E(got[i].cpu().numpy(), want[i], f"thread {i}")


def test_reference_nan_semantics_switch(fu, golden, dev, monkeypatch):
    """field_utils.py:53-54 on targets that coincide with sources: the reference zeroes the LEAF total, so the value
    depends on max_pts.  Default: only the pair is dropped (documented divergence); with REFERENCE_NAN_SEMANTICS the
    reference's values come back for every chunking, exact zeros included."""
    g = golden("potential_semantics")
    cl = torch.from_numpy(g["self_pc"]).to(dev)
    src, tgt = torch.from_numpy(g["mix_src"]).to(dev), torch.from_numpy(g["mix_tgt"]).to(dev)
    hit = g["mix_hit"]
    clean = np.setdiff1d(np.arange(301), hit)
    default = fu.potential(src, tgt).cpu().numpy()
    check_phi(default[clean], g["mix_phi_max15000"][clean], "default, untouched targets")
    assert np.abs(default[hit]).min() > 0 and np.all(g["mix_phi_max15000"][hit] == 0)        # the divergence
    monkeypatch.setattr(fu, "REFERENCE_NAN_SEMANTICS", True)
    for mp in (100, 64, 15000):
        got = fu.potential(src, tgt, max_pts=mp).cpu().numpy()
        want = g[f"mix_phi_max{mp}"]
        assert np.array_equal(got == 0, want == 0), mp
        check_phi(got, want, f"mixed max_pts={mp}")
        got = fu.potential(cl, cl, max_pts=mp).cpu().numpy()
        want = g[f"self_phi_max{mp}"]
        assert np.array_equal(got == 0, want == 0), mp
        check_phi(got, want, f"self max_pts={mp}")
    got = fu.potential(src, tgt, recursive=False).cpu().numpy()
    check_phi(got, g["mix_phi_nonrecursive"], "non-recursive")
    assert np.array_equal(fu.potential(src.cpu(), tgt.cpu(), max_pts=64).numpy() == 0, g["mix_phi_max64"] == 0)
    # inputs without coincidences are untouched by the switch
    a, b = cloud(500, 1).to(dev), cloud(400, 2).to(dev)
    monkeypatch.setattr(fu, "REFERENCE_NAN_SEMANTICS", False)
    base = fu.potential(a, b)
    monkeypatch.setattr(fu, "REFERENCE_NAN_SEMANTICS", True)
    assert torch.equal(fu.potential(a, b), base)


def test_nonfinite_components_are_zeroed_and_reported_on_request(fu, dev, capsys, monkeypatch):
    """field_utils.py:110-115: Inf / NaN totals are zeroed; the reference prints their count, here on request."""
    src = cloud(3000, 71)
    src[5, 3] = float("nan")                       # a NaN moment poisons every target it touches
    tgt = cloud(500, 72, cols=3)
    for ring_mode in ("0", "2"):                   # shared-memory kernel and the window kernel (fused epilogue)
        monkeypatch.setenv("DNP_RING", ring_mode)
        monkeypatch.setattr(fu, "REPORT_NONFINITE", False)
        E = fu.field_grad(src.to(dev), tgt.to(dev))
        assert bool((E == 0).all()) and capsys.readouterr().out == ""
        monkeypatch.setattr(fu, "REPORT_NONFINITE", True)
        fu.field_grad(src.to(dev), tgt.to(dev))
        assert "warning: 1500 inf / nan in field_grad" in capsys.readouterr().out
        fu.field_grad(cloud(100, 73).to(dev), tgt.to(dev))
        assert capsys.readouterr().out == ""


# ------------------------------------------------------------------ a3 lattice sign test
def test_lattice_golden(fu, golden, orc, dev):
    g = golden("lattice")
    sp = torch.from_numpy(g["sphere_pc"]).to(dev)
    m = fu.measure_mean_potential(sp)
    assert m.dim() == 0 and m.device == sp.device
    assert abs(m.item() - float(g["sphere_mean64"])) <= 1e-4 * abs(float(g["sphere_mean64"]))
    assert (m < 0) == (g["sphere_mean"] < 0)
    flipped = sp.clone()
    flipped[:, 3:] *= -1
    assert fu.measure_mean_potential(flipped).item() == -m.item()          # exact antisymmetry
    pp = golden("points_prop")
    pts = torch.from_numpy(np.concatenate([pp["pc_in"][:, :3], pp["full_normals"]], 1)).to(dev)
    m = fu.measure_mean_potential(pts).item()
    want = float(pp["full_mean_potential"])
    assert abs(m - want) <= 1e-4 * abs(want) and (m < 0) == (want < 0)
    assert abs(fu.measure_mean_potential(pts.cpu()).item() - want) <= 1e-4 * abs(want)    # CPU tensor in


def test_lattice_node_potentials(golden, orc, dev):
    import dnp_native
    g = golden("lattice")
    sp = torch.from_numpy(g["sphere_pc"]).to(dev)
    mean, phi = dnp_native.lattice_mean_potential(sp, want_phi=True)
    check_phi(phi.cpu().numpy(), orc.potential(g["sphere_pc"], g["grid"]))
    assert abs(mean.item() - phi.double().mean().item()) <= 1e-6 * abs(mean.item())


@pytest.mark.parametrize("S", [1, 255, 257, 4099, 300001])
@pytest.mark.parametrize("n", [10, 7, 13])
def test_lattice_mean_kernel_vs_float64_oracle(orc, dev, S, n):
    """The mean-only kernel (dnp_lattice.cu: sources in registers, nodes uniform, one running sum per thread)
    against the float64 oracle over the same n^3 nodes; ragged source counts, other lattice sides, a strided
    view, and sources placed exactly ON lattice nodes (coincident pairs contribute 0)."""
    import dnp_native
    pc = cloud(S, 900 + S)
    pc[:, :3] *= 1.6                                         # spans several lattice cells
    k = min(S, 5)
    idx = torch.arange(k)
    pc[:k, 0] = ((idx % n).float() / n - 0.5) * 2            # exactly on nodes
    pc[:k, 1] = (((idx + 1) % n).float() / n - 0.5) * 2
    pc[:k, 2] = (((idx + 2) % n).float() / n - 0.5) * 2
    ar = torch.arange(n ** 3)
    grid = ((torch.stack([ar // (n * n), (ar // n) % n, ar % n], 1).float() / n) - 0.5) * 2     # util.py:26-36
    phi64 = orc.potential(pc.numpy(), grid.numpy(), drop_coincident=True)      # the product's documented rule
    want = float(np.mean(phi64))
    scale = float(np.mean(np.abs(phi64)))
    got = dnp_native.lattice_mean_potential(pc.to(dev), n=n).item()
    assert abs(got - want) <= 1e-4 * abs(want) + 1e-5 * scale, (got, want)
    wide = torch.zeros(S, 9)
    wide[:, 1:7] = pc
    got_view = dnp_native.lattice_mean_potential(wide.to(dev)[:, 1:7], n=n).item()
    assert got_view == got                                    # same kernel, same order: bit-equal
    mean_generic, phi = dnp_native.lattice_mean_potential(pc.to(dev), n=n, want_phi=True)
    assert abs(got - mean_generic.item()) <= 1e-4 * abs(want) + 1e-5 * scale


# ------------------------------------------------------------------ a4 patch-pair flip scores
def test_edge_scores_golden(fu, golden, dev):
    g = golden("edge")
    for k in range(int(g["n_cases"])):
        S, T = torch.from_numpy(g[f"e{k}_S"]).to(dev), torch.from_numpy(g[f"e{k}_T"]).to(dev)
        w, invw = fu.field_edge_calculator(S, T)
        assert isinstance(w, np.ndarray) and w.shape == () and w.dtype == np.float32       # reference's return types
        assert isinstance(invw, np.float32) and invw == -w
        want = float(g[f"e{k}_w64"])
        assert abs(float(w) - want) <= 1e-4 * abs(want), k
        assert fu.field_edge_calculator_bool(S, T) == tuple(g[f"e{k}_bool"])
        assert fu.field_edge_calculator_count(S, T) == tuple(g[f"e{k}_count"])
        w_cpu, _ = fu.field_edge_calculator(S.cpu(), T.cpu())                            # CPU tensors in
        assert abs(float(w_cpu) - want) <= 1e-4 * abs(want)
    cl = torch.from_numpy(g["selfall_pc"]).to(dev)
    want = float(g["selfall_w"])
    assert abs(float(fu.self_interaction_all(cl)) - want) <= 1e-4 * abs(want)
    torch.manual_seed(0)
    np.random.seed(0)
    assert np.isfinite(fu.self_interaction(cl)) and np.isfinite(fu.random_self_interaction(cl))


def test_edge_scores_batched(fu, golden, orc, dev):
    """K3: every edge of a patch graph in one launch equals the per-edge calls."""
    g = golden("patch_prop")
    pc = torch.from_numpy(g["pc_in"]).to(dev)
    perm, off = g["perm"], g["offsets"]
    patches = [torch.from_numpy(perm[off[i]:off[i + 1]]) for i in range(12)]
    edges = [(i, j) for i in range(12) for j in range(12) if i != j][::5]
    w, invw = fu.field_edge_calculator_batched(pc, patches, edges)
    assert w.shape == (len(edges),) and np.array_equal(invw, -w)
    for e, (i, j) in enumerate(edges):
        want, _ = orc.field_edge_calculator(g["pc_in"][patches[i].numpy()], g["pc_in"][patches[j].numpy()])
        assert abs(w[e] - want) <= 1e-4 * abs(want) + 1e-3, (i, j)
        single, _ = fu.field_edge_calculator(pc[patches[i].to(dev)], pc[patches[j].to(dev)])
        assert abs(w[e] - float(single)) <= 1e-4 * abs(want) + 1e-3


# ------------------------------------------------------------------ a5 orientation transfer
def test_reference_field_golden(fu, golden, dev):
    g = golden("transfer")
    ref = torch.from_numpy(g["ref"]).to(dev)
    t3 = torch.from_numpy(g["tgt3"]).to(dev)
    out3 = fu.reference_field(ref, t3)
    assert out3.shape == (700, 6) and out3 is not t3
    assert torch.equal(out3[:, :3], t3)
    assert np.abs(out3[:, 3:].cpu().numpy() - g["out3"][:, 3:]).max() < 1e-5
    t6 = torch.from_numpy(g["tgt6"].copy()).to(dev)
    out6 = fu.reference_field(ref, t6)
    assert out6 is t6                                                      # in place, same object
    assert np.array_equal(out6.cpu().numpy(), g["out6"])                  # every sign identical
    t6c = torch.from_numpy(g["tgt6"].copy())                              # CPU tensor: mutated in place too
    assert fu.reference_field(ref.cpu(), t6c) is t6c
    assert np.array_equal(t6c.numpy(), g["out6"])


# ------------------------------------------------------------------ size-independent properties at BASELINE sizes
def test_properties_at_full_size(fu, dev):
    """1,048,576 x 1,048,576 pairs (BASELINE config 5): linearity in the sources, exact antisymmetry in the
    moments, exact equivariance under a permutation of the targets, and E = -grad(phi) at eps = 0."""
    n = 1 << 20
    src, tgt = cloud(n, 11).to(dev), cloud(n, 12, cols=3).to(dev)
    E = fu.field_grad(src, tgt)
    assert torch.isfinite(E).all()
    # linearity: the two source halves add up (different summation order => FP32 tolerance)
    half = fu.field_grad(src[: n // 2], tgt) + fu.field_grad(src[n // 2:], tgt)
    err = (E - half).norm(dim=-1)
    mag = E.norm(dim=-1)
    assert bool((err <= RTOL * mag + ATOL_E * mag.max()).all())
    # antisymmetry: flipping every moment negates the field bit for bit
    neg = src.clone()
    neg[:, 3:] *= -1
    assert torch.equal(fu.field_grad(neg, tgt), -E)
    # a permutation of the targets permutes the rows, bit for bit
    perm = torch.randperm(n, generator=torch.Generator().manual_seed(5)).to(dev)
    assert torch.equal(fu.field_grad(src, tgt[perm]), E[perm])
    # checksum of checksums: potential of all sources = sum of the potentials of 4 quarters
    phi = fu.potential(src, tgt)
    q = sum(fu.potential(src[i * n // 4:(i + 1) * n // 4], tgt) for i in range(4))
    assert bool(((phi - q).abs() <= RTOL * phi.abs() + ATOL_PHI * phi.abs().max()).all())


def test_field_is_minus_gradient_of_potential(fu, dev):
    """field_grad = -grad_q(potential) at eps = 0 (the reference's sign convention: R = source - target)."""
    src = cloud(4096, 21).to(dev)
    q = (cloud(64, 22, cols=3) * 3.0 + 2.0).double()            # well outside the cloud: smooth field
    h = 1e-2
    E = fu.field_grad(src, q.float().to(dev), eps=0.0).cpu().double()
    grad = torch.zeros_like(E)
    for c in range(3):
        dq = torch.zeros(3, dtype=torch.float64)
        dq[c] = h
        qp, qm = (q + dq).float(), (q - dq).float()
        step = (qp - qm).double()[:, c]
        grad[:, c] = (fu.potential(src, qp.to(dev)).cpu().double() - fu.potential(src, qm.to(dev)).cpu().double()) / step
    assert ((E + grad).norm(dim=-1) <= 2e-2 * E.norm(dim=-1) + 1e-3 * E.norm(dim=-1).max()).all()
