"""GPU parity of the constant-WINDOW all-pairs path (dnp_upairs.cu): sources in uniform registers, fed from a
device-resident constant window, balanced (tile, source) shares per launch.  This is the kernel every field_grad /
potential call above ~1.6e7 pairs takes (field_utils.py:61-116, :12-55), including the benchmarked 1M x 1M size.

Stated FP32 tolerance, per target, against the float64 oracle:
    |E_gpu - E_ref|     <= 1e-4 |E_ref|   + 1e-6 max_m |E_ref|
    |phi_gpu - phi_ref| <= 1e-4 |phi_ref| + 1e-5 max_m |phi_ref|
and the error against float64 may be no worse than 4x the reference's own float32 error on the same inputs
(oracle/torch_port.py is the reference's float32 op chain, bit-equal to field_utils.py).
DNP_RING=2 forces the path at test sizes, =0 disables it, =1 (default) applies the size rule."""
import threading

import numpy as np
import pytest
import torch

from test_gpu_leaf import check_E, check_phi, cloud

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fu():
    import field_utils
    return field_utils


@pytest.fixture()
def ring(monkeypatch):
    def _set(mode, tp=None, fill=None, shape=None, lanes=None):
        """shape: 1 = equal (tile, source) shares per CTA, 2 = whole target tiles per CTA (2 or 4 lanes)"""
        monkeypatch.setenv("DNP_RING", str(mode))
        for name, v in (("DNP_RING_TP", tp), ("DNP_RING_FILL", fill), ("DNP_RING_MODE", shape), ("DNP_RING_LANES", lanes)):
            if v is None:
                monkeypatch.delenv(name, raising=False)
            else:
                monkeypatch.setenv(name, str(v))
    return _set


def rel_err_rows(E, ref):
    E = np.asarray(E, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    return np.linalg.norm(E - ref, axis=-1) / np.maximum(np.linalg.norm(ref, axis=-1), 1e-300)


# ragged sizes around every boundary of the plan: one quad, one fold (224 / 336), one slot (896 / 1344 sources),
# 256- and 1024-target tiles, several chunks over two lanes
SIZES = [(1, 1), (7, 5), (223, 1023), (224, 1024), (225, 1025), (895, 255), (896, 2048), (897, 1025),
         (1344, 300), (1345, 1025), (2689, 2500), (6000, 9001), (3000, 20000)]


@pytest.mark.parametrize("S,M", SIZES)
@pytest.mark.parametrize("tp", [None, 1, 4])
def test_forced_sizes_vs_oracle(fu, orc, dev, ring, S, M, tp):
    ring(2, tp=tp)
    src, tgt = cloud(S, 3000 + S), cloud(M, 4000 + M)
    E = fu.field_grad(src.to(dev), tgt.to(dev)).cpu().numpy()
    phi = fu.potential(src.to(dev), tgt.to(dev)).cpu().numpy()
    check_E(E, orc.field_grad(src.numpy(), tgt.numpy()), f"window {S}x{M} tp={tp}")
    check_phi(phi, orc.potential(src.numpy(), tgt.numpy()), f"window {S}x{M} tp={tp}")


@pytest.mark.parametrize("S,M", SIZES)
@pytest.mark.parametrize("lanes", [2, 4])
def test_whole_tile_launches_vs_oracle(fu, orc, dev, ring, S, M, lanes):
    """The launch shape of the benchmarked sizes (one 1024-target tile per CTA, 2 or 4 lanes), forced at test sizes."""
    ring(2, shape=2, lanes=lanes)
    src, tgt = cloud(S, 5000 + S), cloud(M, 6000 + M)
    E = fu.field_grad(src.to(dev), tgt.to(dev)).cpu().numpy()
    phi = fu.potential(src.to(dev), tgt.to(dev)).cpu().numpy()
    check_E(E, orc.field_grad(src.numpy(), tgt.numpy()), f"whole tiles {S}x{M} lanes={lanes}")
    check_phi(phi, orc.potential(src.numpy(), tgt.numpy()), f"whole tiles {S}x{M} lanes={lanes}")


@pytest.mark.parametrize("fill", [0, 1, 2])
def test_window_fill_modes_agree_bitwise(fu, dev, ring, fill):
    """cudaMemcpyToSymbolAsync (default), a plain device-to-device copy to the window's address and a fill
    kernel must all be seen by the next launch: many chunks, both slots, results bit-identical."""
    src, tgt = cloud(20000, 11).to(dev), cloud(30000, 12, cols=3).to(dev)
    for shape, lanes in ((1, None), (2, 2), (2, 4)):
        ring(2, fill=0, shape=shape, lanes=lanes)
        want = fu.field_grad(src, tgt)
        ring(2, fill=fill, shape=shape, lanes=lanes)
        for _ in range(3):
            assert torch.equal(fu.field_grad(src, tgt), want)


@pytest.mark.parametrize("shape", [1, 2])
def test_coincident_pairs_and_eps(fu, orc, dev, ring, shape):
    """Exact duplicates of targets among the sources (field_utils.py:98-106: such a pair contributes 0).  The
    whole-tile kernel runs a clamp-free fast fold and repairs a fold that met a duplicate (NaN in its partial sums)
    by re-running it with the clamped arithmetic: every target of this test takes that repair path once, and a
    second cloud has duplicates in only a few folds."""
    ring(2, shape=shape)
    src = cloud(3000, 41)
    tgt = src.clone()[:2000]                      # every target coincides with one source
    for eps in (1e-5, 1e-6, 0.0):
        E = fu.field_grad(src.to(dev), tgt.to(dev), eps=eps).cpu().numpy()
        assert np.isfinite(E).all()
        check_E(E, orc.field_grad(src.numpy(), tgt.numpy(), eps=eps), f"coincident eps={eps}")
        check_phi(fu.potential(src.to(dev), tgt.to(dev)).cpu().numpy(),
                  orc.potential(src.numpy(), tgt.numpy(), drop_coincident=True), "coincident potential")
    few = cloud(5000, 43)
    tgt2 = cloud(3000, 44)
    tgt2[[7, 1500, 2999], :3] = few[[4999, 0, 2600], :3]
    E = fu.field_grad(few.to(dev), tgt2.to(dev)).cpu().numpy()
    check_E(E, orc.field_grad(few.numpy(), tgt2.numpy()), "a few duplicates")
    one = cloud(1, 5).to(dev)
    assert bool((fu.field_grad(one, one) == 0).all())            # exactly zero, field_utils.py:98-106
    assert bool((fu.potential(one, one) == 0).all())


def test_size_rule_selects_the_window_path(fu, dev, ring):
    import dnp_native as dn
    ring(1)
    src, tgt = cloud(4000, 51).to(dev), cloud(65536, 52, cols=3).to(dev)     # 2.6e8 pairs: window path
    n0 = dn.launch_count()
    fu.field_grad(src, tgt)
    # pack + ceil(4000 / 896) = 5 chunk launches + reduction
    assert dn.launch_count() - n0 == 1 + 5 + 1
    n0 = dn.launch_count()
    fu.field_grad(src[:500], tgt[:4096])                                       # 2e6 pairs: shared-memory kernel
    assert dn.launch_count() - n0 <= 3
    n0 = dn.launch_count()
    fu.field_grad(src, tgt[:16384])                                            # too few targets per launch
    assert dn.launch_count() - n0 <= 3


@pytest.mark.parametrize("n_src", [700, 1700])
def test_epilogues_gathers_and_masks(fu, orc, dev, ring, n_src):
    """Every fused epilogue + src_idx / tgt_idx / tgt_skip through the window path equals the shared-memory path.
    700 sources = one launch (epilogue applied inside the kernel, cut tiles summed by their last CTA);
    1700 = two launches + the reduction pass."""
    import dnp_native as dn
    g = torch.Generator().manual_seed(61)
    pts = cloud(5000, 62).to(dev)
    src_idx = torch.randperm(5000, generator=g)[:n_src].to(dev)
    tgt_idx = torch.randperm(5000, generator=g)[:2900].to(dev)
    skip = (torch.rand(5000, generator=g) < 0.3).to(dev)
    results = {}
    for mode in (0, 2):
        ring(mode)
        out = {}
        E = torch.zeros(5000, 3, device=dev)
        dn.field_apply(pts, pts, 1e-5, dn.EPI_STORE, out=E, src_idx=src_idx, tgt_idx=tgt_idx, tgt_skip=skip)
        out["store"] = E.clone()
        dn.field_apply(pts, pts, 1e-5, dn.EPI_ACCUM, out=E, src_idx=src_idx, tgt_idx=tgt_idx)
        out["accum"] = E.clone()
        for name, epi in (("ge", dn.EPI_SIGN_GE), ("gt", dn.EPI_SIGN_GT)):
            work = pts.clone()
            dn.field_apply(pts[:, :6].clone(), work, 1e-5, epi, normals=work[:, 3:], src_idx=src_idx, tgt_idx=tgt_idx)
            out[name] = work
        nrm = torch.zeros(5000, 3, device=dev)
        dn.field_apply(pts, pts, 1e-5, dn.EPI_NORMALIZE, out=nrm, src_idx=src_idx)
        out["normalize"] = nrm
        dots = torch.zeros(5000, device=dev)
        dn.field_apply(pts, pts, 1e-5, dn.EPI_DOT, out=dots, ld_out=1, normals=pts[:, 3:], src_idx=src_idx)
        out["dot"] = dots
        results[mode] = out
    a, b = results[0], results[2]
    for k in ("store", "accum", "normalize"):
        err = (a[k] - b[k]).norm(dim=-1)
        mag = a[k].norm(dim=-1)
        assert bool((err <= 1e-4 * mag + 1e-6 * mag.max()).all()), k
    assert bool(((a["dot"] - b["dot"]).abs() <= 1e-4 * a["dot"].abs() + 1e-6 * a["dot"].abs().max()).all())
    for k in ("ge", "gt"):                       # sign decisions: identical except where E.n is within rounding of 0
        diff = (a[k][:, 3:] != b[k][:, 3:]).any(dim=-1)
        assert int(diff.sum()) <= 2, k
    sel = torch.zeros(5000, dtype=torch.bool, device=dev)
    sel[tgt_idx] = True
    assert bool((b["store"][~sel | skip] == 0).all())


def test_host_entry_points_take_the_window_path(fu, orc, ring):
    """CPU tensors in, CPU tensor out (dnp_field_grad_host_f32 / dnp_potential_host_f32): upload, window path,
    download; strided host rows included."""
    ring(2)
    src, tgt = cloud(4000, 71), cloud(3000, 72)
    E = fu.field_grad(src, tgt[:, :3])
    assert not E.is_cuda
    check_E(E.numpy(), orc.field_grad(src.numpy(), tgt.numpy()), "host entry")
    wide = torch.cat([src, torch.zeros(4000, 2)], 1)[:, :6]                 # ld_s = 8
    check_E(fu.field_grad(wide, tgt[:, :3]).numpy(), orc.field_grad(src.numpy(), tgt.numpy()), "host strided")
    check_phi(fu.potential(src, tgt).numpy(), orc.potential(src.numpy(), tgt.numpy()), "host potential")


def test_deterministic_and_reentrant(fu, orc, dev, ring):
    """The window is one per device: concurrent callers on their own streams are queued on it and each gets
    its own answer."""
    ring(2)
    src, tgt = cloud(7000, 81).to(dev), cloud(5000, 82).to(dev)
    assert torch.equal(fu.field_grad(src, tgt), fu.field_grad(src, tgt))
    jobs = [(cloud(2600 + 130 * i, i), cloud(1500 + 70 * i, 100 + i)) for i in range(6)]
    want = [orc.field_grad(s.numpy(), t.numpy()) for s, t in jobs]
    got = [None] * len(jobs)

    def run(i):
        stream = torch.cuda.Stream(device=dev)
        with torch.cuda.stream(stream):
            for _ in range(6):
                got[i] = fu.field_grad(jobs[i][0].to(dev), jobs[i][1].to(dev))
            stream.synchronize()

    threads = [threading.Thread(target=run, args=(i,)) for i in range(len(jobs))]
    [t.start() for t in threads]
    [t.join() for t in threads]
    for i in range(len(jobs)):
        check_E(got[i].cpu().numpy(), want[i], f"thread {i}")


def test_call_is_asynchronous_and_captures_into_a_cuda_graph(fu, orc, dev, ring):
    """No host synchronisation anywhere on the path: the multi-chunk, two-lane call records into a CUDA graph
    (window fills as memcpy nodes) and the replay reproduces the eager result bit for bit."""
    import dnp_native as dn
    ring(2)
    src, tgt = cloud(3000, 91).to(dev), cloud(2000, 92, cols=3).to(dev)
    E = torch.empty(2000, 3, device=dev)
    dn.field_grad(src, tgt, out=E)
    want = E.clone()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    n0 = dn.launch_count()
    with torch.cuda.graph(graph):
        dn.field_grad(src, tgt, out=E)
    assert dn.launch_count() - n0 == 1 + 4 + 1            # pack + ceil(3000 / 896) window launches + reduction
    for _ in range(3):
        E.zero_()
        graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(E, want)
    check_E(E.cpu().numpy(), orc.field_grad(src.cpu().numpy(), tgt.cpu().numpy()), "graph replay")


@pytest.mark.parametrize("S,M", [(500, 9001), (896, 300000), (640, 40000)])
def test_single_launch_calls_fused_fixup(fu, orc, dev, ring, monkeypatch, S, M):
    """One-launch calls: cut tiles are summed in way order by whichever CTA arrives last, so the result is
    bit-identical to the unfused path (zeroed rows + reduction pass), run after run, also through a CUDA graph
    (the arrival counters clean themselves up)."""
    import dnp_native as dn
    src, tgt = cloud(S, 21).to(dev), cloud(M, 22, cols=3).to(dev)
    ring(2)
    n0 = dn.launch_count()
    E = fu.field_grad(src, tgt)
    assert dn.launch_count() - n0 == 2                      # pack into the window + ONE kernel
    for _ in range(3):
        assert torch.equal(fu.field_grad(src, tgt), E)
    phi = fu.potential(src, tgt)
    monkeypatch.setenv("DNP_RING_FUSED", "0")
    n0 = dn.launch_count()
    assert torch.equal(fu.field_grad(src, tgt), E)
    assert dn.launch_count() - n0 == 3                      # pack, kernel, reduction pass
    assert torch.equal(fu.potential(src, tgt), phi)
    monkeypatch.delenv("DNP_RING_FUSED")
    if M <= 40000:
        check_E(E.cpu().numpy(), orc.field_grad(src.cpu().numpy(), tgt.cpu().numpy()), "fused")
    out = torch.empty(M, 3, device=dev)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        dn.field_grad(src, tgt, out=out)
    for _ in range(3):
        out.zero_()
        graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(out, E)


# ------------------------------------------------------------------ the benchmarked sizes, unforced
def test_mid_size_call_in_full_vs_float64_oracle(fu, orc, dev, ring):
    """20 000 x 140 000 in full (2.8e9 pairs, 23 chunks over two lanes, near pairs included): stated tolerance and
    'no worse than 4x the reference's own float32 error'."""
    from oracle import torch_port
    ring(1)
    S, M = 20000, 140000
    g = torch.Generator().manual_seed(2024)
    d = torch.randn(S + M, 3, generator=g)
    d = d / d.norm(dim=-1, keepdim=True)
    xyz = d * 0.45 + torch.randn(S + M, 3, generator=g) * 0.002          # noisy sphere
    pts = torch.cat([xyz, d], 1).float()
    src, tgt = pts[:S].contiguous(), pts[S:].contiguous()
    ref64 = orc.field_grad(src.numpy(), tgt.numpy())
    E = fu.field_grad(src.to(dev), tgt.to(dev)).cpu().numpy()
    check_E(E, ref64, "20000 x 140000 field")
    phi = fu.potential(src.to(dev), tgt.to(dev)).cpu().numpy()
    check_phi(phi, orc.potential(src.numpy(), tgt.numpy()), "20000 x 140000 potential")
    # the reference's own float32 error on a 4096-target subsample of the same problem
    sub = torch.randperm(M, generator=g)[:4096]
    ref32 = torch_port.field_grad(src, tgt[sub]).numpy()
    own = rel_err_rows(ref32, ref64[sub.numpy()])
    ours = rel_err_rows(E[sub.numpy()], ref64[sub.numpy()])
    assert np.median(ours) <= 4 * np.median(own) and ours.max() <= 4 * own.max(), (np.median(ours), ours.max(),
                                                                                   np.median(own), own.max())


def test_benchmark_size_sources_on_a_target_subsample(fu, orc, dev, ring):
    """All 1 048 576 sources of the benchmark (1171 window launches, both lanes) against a 4096-target random
    subsample of its targets; the subsample rows of the FULL 1M x 1M result must equal the subsample call's
    rows to rounding (different share boundaries, same sources)."""
    ring(1)
    S, M = 1 << 20, 1 << 20
    src = cloud(S, 1234)
    tgt = cloud(M, 1235, cols=3)
    g = torch.Generator().manual_seed(7)
    sub = torch.randperm(M, generator=g)[:4096]
    ref64 = orc.field_grad(src.numpy(), tgt[sub].numpy())
    d_src, d_tgt = src.to(dev), tgt.to(dev)
    E_full = fu.field_grad(d_src, d_tgt)
    E_sub = E_full[sub.to(dev)].cpu().numpy()
    check_E(E_sub, ref64, "1M sources x subsample of the full result")
    phi_full = fu.potential(d_src, d_tgt)
    check_phi(phi_full[sub.to(dev)].cpu().numpy(), orc.potential(src.numpy(), tgt[sub].numpy()), "1M potential")
    E_small = fu.field_grad(d_src, d_tgt[sub.to(dev)].contiguous()).cpu().numpy()    # 4.3e9 pairs: window path, TP1
    check_E(E_small, ref64, "1M sources x 4096 targets")
