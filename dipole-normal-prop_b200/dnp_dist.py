"""Multi-GPU layer of the dipole hot path: one process per GPU,
`torch.distributed` (NCCL over NVLink on the B200 box) as plumbing.

The reference is single-process, single-device; this layer is net-new.  Partitioning follows
BASELINE.json's north_star:

  * all-pairs potential / field: TARGETS are block-sharded over the ranks, sources are replicated.
    Every target is independent, so the data path has NO collective; an all_gather of the result is
    optional (`gather=True`) for callers that want the full array on every rank.
  * patch-pair flip scores: every rank reduces ITS slice of each target patch into a per-edge partial
    sum (one launch), then ONE all_reduce(SUM) over the [n_edges] vector combines them.
  * lattice sign test: 1000 nodes are too few to shard; SOURCES are sharded instead (the potential is
    linear in the sources) and the [n^3] node potentials are combined with one all_reduce(SUM).
  * the greedy drivers (a6-a8) are sequential chains: replicas only (e.g. one patch per rank).

`compute` hooks exist so the sharding / collective logic can be exercised on CPU with the gloo backend
(tests/test_dist.py injects the oracle there); the product default is the CUDA library, no fallback.
"""
import torch
import torch.distributed as dist

import dnp_native as _dn


def shard_range(n, rank, world):
    """Block partition of range(n): the first n % world ranks get one extra element."""
    base, rem = divmod(int(n), int(world))
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def _rank_world(group):
    if not dist.is_available() or not dist.is_initialized():
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def _gather_rows(local, n_total, group):
    """all_gather of row blocks produced by shard_range (uneven blocks are padded to the largest)."""
    rank, world = _rank_world(group)
    if world == 1:
        return local
    rows = -(-n_total // world)
    padded = local.new_zeros((rows,) + tuple(local.shape[1:]))
    padded[: local.shape[0]] = local
    parts = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(parts, padded, group=group)
    out = []
    for r in range(world):
        a, b = shard_range(n_total, r, world)
        out.append(parts[r][: b - a])
    return torch.cat(out, dim=0)


def field_grad_sharded(sources, means, eps=1e-5, group=None, gather=False, compute=None):
    """field_grad with the targets block-sharded over the ranks of `group`, sources replicated.
    Returns this rank's [m_local, 3] block (gather=False) or the full [M, 3] array (gather=True)."""
    rank, world = _rank_world(group)
    a, b = shard_range(means.shape[0], rank, world)
    fn = compute or (lambda s, t: _dn.field_grad(s, t, eps=eps))
    local = fn(sources, means[a:b])
    return _gather_rows(local, means.shape[0], group) if gather else local


def potential_sharded(sources, means, group=None, gather=False, compute=None):
    """potential with the targets block-sharded over the ranks, sources replicated."""
    rank, world = _rank_world(group)
    a, b = shard_range(means.shape[0], rank, world)
    fn = compute or (lambda s, t: _dn.potential(s, t))
    local = fn(sources, means[a:b])
    return _gather_rows(local, means.shape[0], group) if gather else local


def measure_mean_potential_sharded(pc, n=10, group=None, compute=None):
    """Lattice sign test with the SOURCES sharded: phi[n^3] partial sums + one all_reduce(SUM)."""
    rank, world = _rank_world(group)
    a, b = shard_range(pc.shape[0], rank, world)
    fn = compute or (lambda s: _dn.lattice_mean_potential(s, n=n, want_phi=True)[1])
    phi = fn(pc[a:b])
    if world > 1:
        dist.all_reduce(phi, op=dist.ReduceOp.SUM, group=group)
    return phi.mean()


def slice_patches(offsets, rank, world):
    """Per-patch block slices of a CSR: returns (gather_index, new_offsets) such that
    perm[gather_index] / new_offsets is the CSR of this rank's share of every patch (shard_range per patch).
    Vectorised on the device of `offsets`: no per-patch Python work."""
    off = torch.as_tensor(offsets, dtype=torch.int64)
    sizes = off[1:] - off[:-1]
    base, rem = sizes // world, sizes % world
    start = rank * base + torch.clamp(rem, max=rank)                    # shard_range(n, rank, world), all patches at once
    count = base + (rank < rem).to(torch.int64)
    new_off = torch.zeros_like(off)
    new_off[1:] = torch.cumsum(count, 0)
    total = int(new_off[-1])
    patch = torch.repeat_interleave(torch.arange(sizes.numel(), device=off.device), count, output_size=total)
    idx = torch.arange(total, device=off.device) - new_off[:-1][patch] + (off[:-1] + start)[patch]
    return idx, new_off


def edge_interactions_sharded(pts, perm, offsets, edges, eps=1e-5, group=None, compute=None):
    """Per-edge interactions sum_t n_t . E_S(t) with the targets of EVERY patch sliced over the ranks
    and the partial sums combined by a single all_reduce(SUM) over the edge vector."""
    rank, world = _rank_world(group)
    gidx, t_off = slice_patches(torch.as_tensor(offsets).to(perm.device), rank, world)
    t_perm = perm[gidx]
    fn = compute or (lambda tp, to: _dn.edge_interactions_batched(pts, perm, offsets, edges, eps=eps,
                                                                   tgt_perm=tp, tgt_offsets=to))
    partial = fn(t_perm, t_off.to(perm.device))
    if world > 1:
        dist.all_reduce(partial, op=dist.ReduceOp.SUM, group=group)
    return partial


def edge_scores_sharded(pts, perm, offsets, edges, eps=1e-5, group=None, compute=None):
    """(w, invw) of field_edge_calculator (field_utils.py:145-160) for an edge list, multi-GPU."""
    inter = edge_interactions_sharded(pts, perm, offsets, edges, eps=eps, group=group, compute=compute)
    sizes = (offsets[1:] - offsets[:-1]).to(device=inter.device, dtype=torch.float32)
    e = torch.as_tensor(edges).to(inter.device)
    w = (inter * 2) / sizes[e[:, 0]] * sizes[e[:, 1]]
    return w, w * -1
