"""The allocation-free entry point dnp_lattice_mean_potential_noalloc_f32 (include/dipole.h: no counterpart in the
reference): its partial sums live in a `static __device__` array of the
library instead of a caller workspace.  Called through the C ABI (ctypes) and compared with the existing entry point
dnp_lattice_mean_potential_f32 on the same inputs -- same kernels, same launch shape, so the results are required to
be within the lattice tests' tolerance (1e-6 relative here; in practice identical bits) -- on two calls in a row that
share the scratch, with no device allocation across the calls.  This file is named to be collected LAST."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _cloud(n, seed, dev):
    g = torch.Generator().manual_seed(seed)
    d = torch.nn.functional.normalize(torch.randn(n, 3, generator=g), dim=1)
    pc = torch.cat([d * 0.45 + 0.01 * torch.randn(n, 3, generator=g), d], 1).float()
    pc[torch.rand(n, generator=g) < 0.3, 3:] *= -1
    return pc.to(dev)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _existing(lib, pc, n, mean, ws, stream):
    rc = lib.dnp_lattice_mean_potential_f32(_ptr(pc), pc.shape[0], pc.stride(0), n, None, _ptr(mean), _ptr(ws),
                                            ws.numel(), stream)
    assert rc == 0, lib.dnp_last_error()


def test_noalloc_lattice_mean_matches_the_workspace_entry_point(dev):
    import dnp_native
    lib = dnp_native.load()
    torch.cuda.set_device(dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    # 600 000 sources take the 8-sources-per-thread plan (444 CTAs), 7 001 the 2-per-thread plan with fewer CTAs: the
    # second call reads fewer scratch slots than the first one wrote
    clouds = [(_cloud(600000, 1, dev), 10), (_cloud(7001, 2, dev), 10), (_cloud(5000, 3, dev), 7), (_cloud(1, 4, dev), 10)]
    ws = torch.empty(lib.dnp_pairs_workspace_bytes(600000, 1000), dtype=torch.uint8, device=dev)
    want = [torch.zeros((), device=dev) for _ in clouds]
    got = [torch.full((), float("nan"), device=dev) for _ in clouds]
    for (pc, n), w in zip(clouds, want):
        _existing(lib, pc, n, w, ws, stream)
    torch.cuda.synchronize(dev)

    def counters():
        return (torch.cuda.memory_reserved(dev), torch.cuda.memory_stats(dev)["allocation.all.allocated"],
                torch.cuda.mem_get_info(dev)[0])

    for _ in range(2):                                        # two rounds in a row over the shared scratch
        before = counters()
        for (pc, n), g in zip(clouds, got):
            rc = lib.dnp_lattice_mean_potential_noalloc_f32(_ptr(pc), pc.shape[0], pc.stride(0), n, _ptr(g), stream)
            assert rc == 0, lib.dnp_last_error()
        torch.cuda.synchronize(dev)
        after = counters()
        # around the calls themselves (the checks below use torch ops, which take blocks from torch's pool): torch
        # allocated nothing, and the device's free memory did not shrink (no cudaMalloc inside the library)
        assert after[0] == before[0] and after[1] == before[1]
        assert after[2] >= before[2] - (2 << 20)
        for g, w, (pc, n) in zip(got, want, clouds):
            assert torch.isfinite(g)
            assert abs(g.item() - w.item()) <= 1e-6 * abs(w.item()) + 1e-12, (pc.shape[0], n, g.item(), w.item())
            g.fill_(float("nan"))
        torch.cuda.synchronize(dev)

    # the Python wrapper, and the empty cloud
    m = dnp_native.lattice_mean_potential_noalloc(clouds[0][0], out=got[0])
    torch.cuda.synchronize(dev)
    assert abs(m.item() - want[0].item()) <= 1e-6 * abs(want[0].item())
    rc = lib.dnp_lattice_mean_potential_noalloc_f32(None, 0, 6, 10, _ptr(got[1]), stream)
    torch.cuda.synchronize(dev)
    assert rc == 0 and got[1].item() == 0.0


def test_noalloc_limits_are_rejected_with_a_message(dev):
    import dnp_native
    lib = dnp_native.load()
    pc = _cloud(64, 5, dev)
    out = torch.zeros((), device=dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    too_many = (1 << 24) + 1
    assert lib.dnp_lattice_mean_potential_noalloc_f32(_ptr(pc), too_many, 6, 10, _ptr(out), stream) == -1
    assert b"at most 16777216" in lib.dnp_last_error()
    assert lib.dnp_lattice_mean_potential_noalloc_f32(_ptr(pc), 64, 6, 256, _ptr(out), stream) == -1
    assert b"lattice side" in lib.dnp_last_error()
    assert lib.dnp_lattice_mean_potential_noalloc_f32(_ptr(pc), 64, 6, 0, _ptr(out), stream) == -1
    assert lib.dnp_lattice_mean_potential_noalloc_f32(_ptr(pc), 64, 5, 10, _ptr(out), stream) == -1
    assert lib.dnp_lattice_mean_potential_noalloc_f32(_ptr(pc), 64, 6, 10, None, stream) == -1
    assert lib.dnp_lattice_mean_potential_noalloc_f32(_ptr(pc), 64, 6, 10, _ptr(out), stream) == 0    # still usable
    torch.cuda.synchronize(dev)
