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


# ---------------------------------------------------------------------------- sharded greedy patch propagation
class _OracleBackend:
    """CPU stand-in for the library calls of dnp_dist's sharded drivers: the reference's op chain
    (oracle/torch_port.py) with the sharded protocol of dnp_patch_propagate_sharded_f32 spelled out step by step --
    this rank adds the chosen patch's field to ITS rows, sums E.n over ITS patches, one all_reduce, every rank takes
    the arg-max."""

    def flattest_patch(self, pts, perm, offsets):
        from oracle import torch_drivers
        groups = [perm[offsets[p]:offsets[p + 1]] for p in range(len(offsets) - 1)]
        return torch_drivers._flattest(pts, groups)

    def greedy_loop(self, pts, perm, offsets, start, E, diffuse, rest_perm, rest_offsets, tgt_rows, owned):
        from oracle.torch_port import field_grad
        P = len(offsets) - 1
        rows = [perm[offsets[p]:offsets[p + 1]] for p in range(P)]
        rest = [rest_perm[rest_offsets[p]:rest_offsets[p + 1]] for p in range(P)] if rest_perm is not None else None
        is_tgt = torch.zeros(pts.shape[0], dtype=torch.bool)
        is_tgt[tgt_rows] = True
        settled = torch.zeros(pts.shape[0], dtype=torch.bool)
        active = torch.ones(P, dtype=torch.bool)
        cur = start
        self.order = [cur]
        for step in range(P):
            active[cur] = False
            settled[rows[cur]] = True
            skip = torch.zeros_like(settled)
            if diffuse:
                skip[rows[cur]] = True                       # :256-261 only the source patch itself
            else:
                skip = settled                               # :264-265
            mine = torch.nonzero(is_tgt & ~skip).flatten()
            if mine.numel():
                E[mine] += field_grad(pts[rows[cur]], pts[mine, :3], eps=1e-5)
            if step == P - 1:
                break
            inter = torch.zeros(P)
            for p in range(P):
                if active[p] and owned[p]:
                    inter[p] = (E[rows[p]] * pts[rows[p], 3:]).sum(dim=-1).sum()
            dist.all_reduce(inter)
            cand = inter.abs() * active
            cur = int(torch.argmax(torch.where(active, cand, torch.full_like(cand, -1.0))))
            self.order.append(cur)
            if inter[cur] < 0:                               # :250-252
                pts[rows[cur], 3:] *= -1
                if rest is not None:
                    pts[rest[cur], 3:] *= -1

    def sign_strict(self, E, pts, rows):
        dots = (E[rows] * pts[rows, 3:]).sum(-1)
        pts[rows, 3:] = pts[rows, 3:] * ((dots > 0).float() * 2 - 1)[:, None]

    def transfer_sign(self, pts, src_rows, tgt_rows):
        from oracle.torch_port import field_grad
        E = field_grad(pts[src_rows], pts[tgt_rows, :3], eps=1e-5)
        dots = (E * pts[tgt_rows, 3:]).sum(-1)
        pts[tgt_rows, 3:] = pts[tgt_rows, 3:] * ((dots > 0).float() * 2 - 1)[:, None]


def _surface(n, seed):
    g = torch.Generator().manual_seed(seed)
    d = torch.nn.functional.normalize(torch.randn(n, 3, generator=g), dim=1)
    pc = torch.cat([d * torch.tensor([0.45, 0.35, 0.25]), torch.nn.functional.normalize(d / torch.tensor([0.45, 0.35, 0.25]), dim=1)], 1)
    pc[torch.rand(n, generator=g) < 0.5, 3:] *= -1
    return pc.float()


def _patches(pc, P):
    order = torch.argsort(pc[:, 0] * 5 + pc[:, 1], stable=True)
    return list(torch.chunk(order, P))


def _prop_worker(rank, world, port, out_dir):
    for p in (ROOT, PKG):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import dnp_dist
    out = {}
    for diffuse in (False, True):
        pc = _surface(900, 4)
        chunks = _patches(pc, 7)                                   # 7 patches on 2 ranks: uneven ownership
        reps = [(c[:40], c[40:]) for c in chunks]
        be = _OracleBackend()
        dnp_dist.strongest_field_propagation_reps_sharded(pc, reps, diffuse=diffuse, backend=be)
        out[f"reps_{diffuse}"] = (pc.clone(), list(be.order))
        pc = _surface(700, 5)
        chunks = _patches(pc, 5)
        be = _OracleBackend()
        dnp_dist.strongest_field_propagation_sharded(pc, [(i, c) for i, c in enumerate(chunks)][:3], chunks,
                                                     diffuse=diffuse, backend=be)
        out[f"all_{diffuse}"] = (pc.clone(), list(be.order))
    assert dnp_dist.patch_owner_mask(7, rank, world).tolist() == [int(p % world == rank) for p in range(7)]
    torch.save(out, os.path.join(out_dir, f"prop_rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_sharded_greedy_propagation_world_size_2_gloo(tmp_path):
    """Patch ownership + one all_reduce per step + the exchanges after the loop: both ranks end with the same cloud,
    and it is the cloud the single-process driver (the reference's op sequence, oracle/torch_drivers.py) produces."""
    from oracle import torch_drivers
    world = 2
    mp.spawn(_prop_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    res = [torch.load(os.path.join(tmp_path, f"prop_rank{r}.pt")) for r in range(world)]
    for diffuse in (False, True):
        for key in (f"reps_{diffuse}", f"all_{diffuse}"):
            assert torch.equal(res[0][key][0], res[1][key][0]), key          # identical on every rank, bit for bit
            assert res[0][key][1] == res[1][key][1], key
        pc = _surface(900, 4)
        chunks = _patches(pc, 7)
        trace = []
        torch_drivers.propagate_representatives(pc, [(c[:40], c[40:]) for c in chunks], diffuse=diffuse, trace=trace)
        assert trace == res[0][f"reps_{diffuse}"][1]
        assert torch.equal(pc, res[0][f"reps_{diffuse}"][0])
        pc = _surface(700, 5)
        chunks = _patches(pc, 5)
        trace = []
        torch_drivers.propagate_patches(pc, [(i, c) for i, c in enumerate(chunks)][:3], chunks, diffuse=diffuse, trace=trace)
        assert trace == res[0][f"all_{diffuse}"][1]
        assert torch.equal(pc, res[0][f"all_{diffuse}"][0])
