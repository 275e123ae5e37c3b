"""Multi-GPU host logic on CPU: world_size-2 `gloo` process group, the oracle injected as the compute hook.
Checks that target sharding + (optional) all_gather, source sharding + all_reduce for the lattice test and
per-patch target slicing + ONE all_reduce for the edge scores reproduce the single-process results."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import PKG, ROOT


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _cloud(n, seed):
    g = torch.Generator().manual_seed(seed)
    xyz = torch.rand(n, 3, generator=g) - 0.5
    nr = torch.randn(n, 3, generator=g)
    return torch.cat([xyz, nr / nr.norm(dim=-1, keepdim=True)], 1).float()


def _worker(rank, world, port, out_dir):
    for p in (ROOT, PKG):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import dnp_dist
    from oracle import oracle as orc

    src, tgt = _cloud(301, 1), _cloud(207, 2)          # uneven shards on purpose
    f_field = lambda s, t: torch.from_numpy(orc.field_grad(s.numpy(), t.numpy()))
    f_phi = lambda s, t: torch.from_numpy(orc.potential(s.numpy(), t.numpy()))
    E_local = dnp_dist.field_grad_sharded(src, tgt, compute=f_field)
    E_full = dnp_dist.field_grad_sharded(src, tgt, gather=True, compute=f_field)
    phi_full = dnp_dist.potential_sharded(src, tgt, gather=True, compute=f_phi)
    a, b = dnp_dist.shard_range(207, rank, world)
    assert E_local.shape == (b - a, 3)

    grid = orc.gen_grid(10)
    f_lat = lambda s: torch.from_numpy(orc.potential(s.numpy(), grid))
    mean = dnp_dist.measure_mean_potential_sharded(src, compute=f_lat)

    # edge scores: 5 patches of uneven size, 6 edges
    perm = torch.randperm(301, generator=torch.Generator().manual_seed(3))
    offsets = torch.tensor([0, 40, 41, 120, 250, 301])
    edges = torch.tensor([[0, 1], [1, 0], [2, 3], [3, 2], [4, 0], [1, 1]])

    def f_edge(t_perm, t_off):
        vals = []
        for s_p, t_p in edges.tolist():
            S = src[perm[offsets[s_p]:offsets[s_p + 1]]].numpy()
            T = src[t_perm[t_off[t_p]:t_off[t_p + 1]]].numpy()
            vals.append(orc.edge_interaction(S, T) if len(T) else 0.0)
        return torch.tensor(vals, dtype=torch.float64)
    w, invw = dnp_dist.edge_scores_sharded(src, perm, offsets, edges, compute=f_edge)

    torch.save({"E_local": E_local, "E_full": E_full, "phi_full": phi_full, "mean": mean, "w": w, "range": (a, b)},
               os.path.join(out_dir, f"rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_world_size_2_gloo(tmp_path, orc):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    res = [torch.load(os.path.join(tmp_path, f"rank{r}.pt")) for r in range(world)]

    src, tgt = _cloud(301, 1), _cloud(207, 2)
    E = orc.field_grad(src.numpy(), tgt.numpy())
    phi = orc.potential(src.numpy(), tgt.numpy())
    covered = []
    for r in range(world):
        a, b = res[r]["range"]
        covered.extend(range(a, b))
        assert np.array_equal(res[r]["E_local"].numpy(), E[a:b])           # a rank's block = the same rows
        assert np.array_equal(res[r]["E_full"].numpy(), E)                 # all_gather reassembles every row
        assert np.array_equal(res[r]["phi_full"].numpy(), phi)
    assert covered == list(range(207))                                       # disjoint, complete cover

    want_mean = orc.measure_mean_potential(src.numpy())
    for r in range(world):
        assert abs(res[r]["mean"].item() - want_mean) <= 1e-12 * abs(want_mean)   # source-sharded + all_reduce

    perm = torch.randperm(301, generator=torch.Generator().manual_seed(3))
    offsets = [0, 40, 41, 120, 250, 301]
    edges = [[0, 1], [1, 0], [2, 3], [3, 2], [4, 0], [1, 1]]
    for e, (s_p, t_p) in enumerate(edges):
        S = src[perm[offsets[s_p]:offsets[s_p + 1]]].numpy()
        T = src[perm[offsets[t_p]:offsets[t_p + 1]]].numpy()
        want, _ = orc.field_edge_calculator(S, T)
        for r in range(world):
            assert abs(res[r]["w"][e].item() - want) <= 1e-9 * abs(want) + 1e-12, (e, r)
    assert torch.equal(res[0]["w"], res[1]["w"])                             # all_reduce: same on every rank


def test_shard_range_properties():
    import dnp_dist
    for n in (0, 1, 7, 8, 1000, 1 << 20):
        for world in (1, 2, 3, 4, 8):
            edges = [dnp_dist.shard_range(n, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in edges]
            assert max(sizes) - min(sizes) <= 1


def test_slice_patches_covers_every_patch():
    import dnp_dist
    offsets = torch.tensor([0, 5, 5, 6, 106])
    seen = []
    for r in range(3):
        idx, off = dnp_dist.slice_patches(offsets, r, 3)
        assert off[0] == 0 and off[-1] == len(idx) and len(off) == 5
        seen.extend(idx.tolist())
    assert sorted(seen) == list(range(106))


def test_single_process_is_identity(orc):
    """Without an initialised process group the sharded calls are the plain calls (world = 1)."""
    import dnp_dist
    src, tgt = _cloud(50, 1), _cloud(20, 2)
    f = lambda s, t: torch.from_numpy(orc.field_grad(s.numpy(), t.numpy()))
    assert torch.equal(dnp_dist.field_grad_sharded(src, tgt, gather=True, compute=f), f(src, tgt))
