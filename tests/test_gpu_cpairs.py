"""GPU parity of the constant-bank ("by value") all-pairs path (dnp_cpairs.cu), the kernel large
field_grad / potential calls take (field_utils.py:61-116, :12-55).  Same stated FP32 tolerance as
tests/test_gpu_leaf.py; DNP_CONST_PATH=2 forces the path at test sizes, =0 disables it."""
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
def path(monkeypatch):
    def _set(mode, lanes=None):
        monkeypatch.setenv("DNP_RING", "0")                 # these tests are about the by-value kernel
        monkeypatch.setenv("DNP_CONST_PATH", str(mode))
        if lanes is None:
            monkeypatch.delenv("DNP_CONST_LANES", raising=False)
        else:
            monkeypatch.setenv("DNP_CONST_LANES", str(lanes))
    return _set


@pytest.mark.parametrize("S,M", [(1, 1), (7, 5), (223, 1023), (224, 1024), (897, 1025), (1281, 1025), (2689, 2500), (6000, 9001)])
def test_forced_sizes_vs_oracle(fu, orc, dev, path, S, M):
    path(2)
    src, tgt = cloud(S, 3000 + S), cloud(M, 4000 + M)
    E = fu.field_grad(src.to(dev), tgt.to(dev)).cpu().numpy()
    phi = fu.potential(src.to(dev), tgt.to(dev)).cpu().numpy()
    check_E(E, orc.field_grad(src.numpy(), tgt.numpy()), f"const {S}x{M}")
    check_phi(phi, orc.potential(src.numpy(), tgt.numpy()), f"const {S}x{M}")


@pytest.mark.parametrize("lanes", [1, 2, 5, 8])
def test_lane_counts(fu, orc, dev, path, lanes):
    path(2, lanes)
    src, tgt = cloud(9000, 31), cloud(2100, 32)
    E = fu.field_grad(src.to(dev), tgt.to(dev)).cpu().numpy()
    check_E(E, orc.field_grad(src.numpy(), tgt.numpy()), f"lanes={lanes}")


def test_coincident_pairs_and_eps(fu, orc, dev, path):
    path(2)
    src = cloud(3000, 41)
    tgt = src.clone()[:2000]                      # every target coincides with one source
    for eps in (1e-5, 1e-6, 0.0):
        E = fu.field_grad(src.to(dev), tgt.to(dev), eps=eps).cpu().numpy()
        assert np.isfinite(E).all()
        check_E(E, orc.field_grad(src.numpy(), tgt.numpy(), eps=eps), f"coincident eps={eps}")


def test_threshold_selects_the_path_and_matches_the_shared_memory_kernel(fu, dev, path):
    import dnp_native as dn
    S, M = 16384, 131072
    src, tgt = cloud(S, 51).to(dev), cloud(M, 52, cols=3).to(dev)
    path(0)
    n0 = dn.launch_count()
    E_smem = fu.field_grad(src, tgt)
    phi_smem = fu.potential(src, tgt)
    smem_launches = dn.launch_count() - n0
    path(1)
    n0 = dn.launch_count()
    E_c = fu.field_grad(src, tgt)
    phi_c = fu.potential(src, tgt)
    const_launches = dn.launch_count() - n0
    # pack + chunk launches (19 blocks of 896 field sources, 13 blocks of 1280 potential sources) + lane reduction
    assert const_launches == (1 + 19 + 1) + (1 + 13 + 1) and smem_launches < const_launches
    err = (E_c - E_smem).norm(dim=-1)
    mag = E_smem.norm(dim=-1)
    assert bool((err <= 1e-4 * mag + 1e-6 * mag.max()).all())
    assert bool(((phi_c - phi_smem).abs() <= 1e-4 * phi_smem.abs() + 1e-5 * phi_smem.abs().max()).all())
    # below the threshold the asynchronous shared-memory kernel runs
    n0 = dn.launch_count()
    fu.field_grad(src[:8192], tgt)
    assert dn.launch_count() - n0 <= 3
    # second branch of the rule: >= 32 768 sources x >= 65 536 targets (37 blocks of 896 sources)
    big_src = cloud(32768, 53).to(dev)
    n0 = dn.launch_count()
    fu.field_grad(big_src, tgt[:65536])
    assert dn.launch_count() - n0 == 1 + 37 + 1
    n0 = dn.launch_count()
    fu.field_grad(big_src, tgt[:65535])
    assert dn.launch_count() - n0 <= 3


def test_epilogues_gathers_and_masks(fu, orc, dev, path):
    """Every fused epilogue + src_idx / tgt_idx / tgt_skip through the by-value path equals the shared-memory path."""
    import dnp_native as dn
    g = torch.Generator().manual_seed(61)
    pts = cloud(5000, 62).to(dev)
    src_idx = torch.randperm(5000, generator=g)[:1700].to(dev)
    tgt_idx = torch.randperm(5000, generator=g)[:2900].to(dev)
    skip = (torch.rand(5000, generator=g) < 0.3).to(dev)
    results = {}
    for mode in (0, 2):
        path(mode)
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
    # untouched rows stay untouched: skipped / unselected targets of the STORE call are still zero
    sel = torch.zeros(5000, dtype=torch.bool, device=dev)
    sel[tgt_idx] = True
    assert bool((b["store"][~sel | skip] == 0).all())


def test_host_entry_passes_sources_by_value(fu, orc, path):
    path(2)
    src, tgt = cloud(4000, 71), cloud(3000, 72)
    E = fu.field_grad(src, tgt[:, :3])                       # CPU tensors in, CPU tensor out
    assert not E.is_cuda
    check_E(E.numpy(), orc.field_grad(src.numpy(), tgt.numpy()), "host by value")
    wide = torch.cat([src, torch.zeros(4000, 2)], 1)[:, :6]  # ld_s = 8: strided host rows take the upload path
    E2 = fu.field_grad(wide, tgt[:, :3])
    check_E(E2.numpy(), orc.field_grad(src.numpy(), tgt.numpy()), "host strided")


def test_deterministic_and_reentrant(fu, orc, dev, path):
    path(2)
    src, tgt = cloud(7000, 81).to(dev), cloud(5000, 82).to(dev)
    assert torch.equal(fu.field_grad(src, tgt), fu.field_grad(src, tgt))
    jobs = [(cloud(2600 + 130 * i, i), cloud(1500 + 70 * i, 100 + i)) for i in range(6)]
    want = [orc.field_grad(s.numpy(), t.numpy()) for s, t in jobs]
    got = [None] * len(jobs)

    def run(i):
        stream = torch.cuda.Stream(device=dev)
        with torch.cuda.stream(stream):
            for _ in range(4):
                got[i] = fu.field_grad(jobs[i][0].to(dev), jobs[i][1].to(dev))
            stream.synchronize()

    threads = [threading.Thread(target=run, args=(i,)) for i in range(len(jobs))]
    [t.start() for t in threads]
    [t.join() for t in threads]
    for i in range(len(jobs)):
        check_E(got[i].cpu().numpy(), want[i], f"thread {i}")


def test_stream_capture_keeps_the_asynchronous_kernel(fu, dev, path):
    """The by-value path synchronises once per call, so a capturing stream gets the shared-memory kernel."""
    path(2)
    src, tgt = cloud(3000, 91).to(dev), cloud(2000, 92, cols=3).to(dev)
    E = torch.empty(2000, 3, device=dev)
    import dnp_native as dn
    dn.field_grad(src, tgt, out=E)
    want = E.clone()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        dn.field_grad(src, tgt, out=E)
    E.zero_()
    graph.replay()
    torch.cuda.synchronize()
    err = (E - want).norm(dim=-1)
    assert bool((err <= 1e-4 * want.norm(dim=-1) + 1e-6 * want.norm(dim=-1).max()).all())
