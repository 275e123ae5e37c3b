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
  * the greedy PATCH drivers (a7, a8) on big clouds: the chain of steps stays sequential, but a step's field update
    and its per-patch sums are sharded by PATCH OWNERSHIP ("inner field updates can be target-sharded,
    the arg-max then needs an all-reduce per step -- worthwhile for N >~ 1e6"): the cloud is replicated, patch p
    belongs to rank p % world, a rank's targets are the rows of its patches.  One all_reduce(SUM) of the [P]
    interaction vector per step (zeros for patches owned elsewhere, so the sum is exact and every rank takes the
    same arg-max and applies the same flips); the final reps -> rest field is sharded over the remaining rows.
    The per-point driver a6 (one 1 x N field per step) stays single-GPU: replicas only.

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


# ---------------------------------------------------------------------------- sharded greedy patch propagation
def patch_owner_mask(n_patches, rank, world, device=None):
    """uint8 [P]: 1 for the patches rank `rank` owns.  Round robin: spatially neighbouring patches (similar sizes,
    visited around the same time) land on different ranks."""
    return (torch.arange(n_patches, device=device) % world == rank).to(torch.uint8)


def owned_rows(index_lists, owned):
    """The rows of the owned patches, concatenated in patch order (this rank's target rows)."""
    mine = [idx for idx, o in zip(index_lists, owned.tolist()) if o]
    if not mine:
        return index_lists[0].new_zeros(0) if index_lists else torch.zeros(0, dtype=torch.int64)
    return torch.cat(mine)


def sum_disjoint(t, group=None):
    """all_reduce(SUM) of a tensor whose non-zero entries are disjoint across the ranks (everything a rank does not
    own is exactly 0): exact, and identical on every rank."""
    if _rank_world(group)[1] > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t


_comm_ready = {"key": None}


def ensure_comm(group=None):
    """The C ABI's communicator (dnp_comm_init) for the ranks of the default group, created once per process: the
    128-byte NCCL id travels through torch.distributed, everything after that is enqueued by the library."""
    rank, world = _rank_world(group)
    key = (rank, world, torch.cuda.current_device())
    if _comm_ready["key"] == key:
        return
    if _comm_ready["key"] is not None:
        _dn.comm_destroy()
    dev = torch.device("cuda", torch.cuda.current_device())
    ident = (_dn.comm_unique_id() if rank == 0 else torch.zeros(128, dtype=torch.uint8)).to(dev)
    if world > 1:
        dist.broadcast(ident, 0, group=group)
    _dn.comm_init(ident.cpu(), rank, world)
    _comm_ready["key"] = key


def release_comm():
    if _comm_ready["key"] is not None:
        _dn.comm_destroy()
        _comm_ready["key"] = None


class _CudaBackend:
    """The library calls the sharded drivers are made of (tests/test_dist.py substitutes the CPU oracle)."""

    def __init__(self, group):
        self.group = group

    def flattest_patch(self, pts, perm, offsets):
        import field_utils
        return field_utils._flattest_patch(pts, perm, offsets)

    def greedy_loop(self, pts, perm, offsets, start, E, diffuse, rest_perm, rest_offsets, tgt_rows, owned):
        if _rank_world(self.group)[1] == 1:          # one rank owns everything: the single-GPU loop, no communicator
            _dn.patch_propagate(pts, perm, offsets, start, E, diffuse=diffuse, eps=1e-5, rest_perm=rest_perm,
                                rest_offsets=rest_offsets, tgt_idx=tgt_rows)
            return
        ensure_comm(self.group)
        _dn.patch_propagate(pts, perm, offsets, start, E, diffuse=diffuse, eps=1e-5, rest_perm=rest_perm,
                            rest_offsets=rest_offsets, tgt_idx=tgt_rows, owned=owned)

    def sign_strict(self, E, pts, rows):
        _dn.field_sign(E, pts, idx=rows, strict=True)

    def transfer_sign(self, pts, src_rows, tgt_rows):
        _dn.field_apply(pts, pts, 1e-5, _dn.EPI_SIGN_GT, normals=pts[:, 3:], src_idx=src_rows, tgt_idx=tgt_rows)


def _csr(index_lists, device):
    sizes = torch.tensor([0] + [int(i.numel()) for i in index_lists], dtype=torch.int64)
    perm = torch.cat(index_lists) if index_lists else torch.zeros(0, dtype=torch.int64, device=device)
    return perm.to(device), torch.cumsum(sizes, 0)


def strongest_field_propagation_reps_sharded(input_pc, reps, diffuse=False, weights=None, group=None, backend=None):
    """field_utils.strongest_field_propagation_reps (field_utils.py:207-282) on the ranks of `group`: every rank passes
    the SAME cloud and patches and ends with the same oriented cloud.  input_pc [N, 6] float32 on this rank's GPU,
    mutated in place; reps = [(rep_idx, rest_idx)] per patch."""
    rank, world = _rank_world(group)
    be = backend or _CudaBackend(group)
    with torch.no_grad():
        pts = input_pc
        dev = pts.device
        N = pts.shape[0]
        if weights is not None:
            weights = weights.to(device=dev, dtype=torch.float32).clamp(0.1, 1)          # :212-215
            pts[:, 3:] = pts[:, 3:] * weights[:, None]
        rep_idx = [torch.as_tensor(rep, dtype=torch.int64, device=dev) for rep, _ in reps]
        rest_idx = [torch.as_tensor(rest, dtype=torch.int64, device=dev) for _, rest in reps]
        P = len(rep_idx)
        perm, offsets = _csr(rep_idx, dev)
        rest_perm, rest_offsets = _csr(rest_idx, dev)
        E = torch.zeros(N, 3, dtype=torch.float32, device=dev)
        oriented = torch.zeros(N, dtype=torch.bool, device=dev)
        oriented[perm] = True
        if P > 0:
            start = be.flattest_patch(pts, perm, offsets)                                 # :230-233, same on all ranks
            owned = patch_owner_mask(P, rank, world, device=dev)
            be.greedy_loop(pts, perm, offsets, start, E, diffuse, rest_perm, rest_offsets,
                           owned_rows(rep_idx, owned), owned)                            # :239-265
        if diffuse:                                                                       # :267-271
            sum_disjoint(E, group)             # every rank filled the rows of its own patches only
            be.sign_strict(E, pts, perm)
        rest_rows = torch.nonzero(~oriented).flatten()                                    # :273-276
        if rest_rows.numel() > 0:
            a, b = shard_range(rest_rows.numel(), rank, world)
            mine = rest_rows[a:b]
            if mine.numel() > 0:
                be.transfer_sign(pts, torch.nonzero(oriented).flatten(), mine)
            new_normals = torch.zeros(rest_rows.numel(), 3, dtype=torch.float32, device=dev)
            new_normals[a:b] = pts[mine, 3:]
            sum_disjoint(new_normals, group)
            pts[rest_rows, 3:] = new_normals
        if weights is not None:                                                           # :280-282
            pts[:, 3:] = pts[:, 3:] / weights[:, None]
    return None


def strongest_field_propagation_sharded(pts, patches, all_patches, diffuse=False, weights=None, group=None,
                                        backend=None):
    """field_utils.strongest_field_propagation (field_utils.py:286-348) on the ranks of `group` (same contract as the
    reps variant: replicated inputs, identical result on every rank)."""
    rank, world = _rank_world(group)
    be = backend or _CudaBackend(group)
    with torch.no_grad():
        dev = pts.device
        N = pts.shape[0]
        if weights is not None:
            weights = weights.to(device=dev, dtype=torch.float32).clamp(0.1, 1)          # :288-291
            pts[:, 3:] = pts[:, 3:] * weights[:, None]
        idx = [torch.as_tensor(i, dtype=torch.int64, device=dev) for i in all_patches]
        P = len(idx)
        perm, offsets = _csr(idx, dev)
        E = torch.zeros(N, 3, dtype=torch.float32, device=dev)
        if P > 0:
            start = be.flattest_patch(pts, perm, offsets)                                 # :303-306
            owned = patch_owner_mask(P, rank, world, device=dev)
            be.greedy_loop(pts, perm, offsets, start, E, diffuse, None, None, owned_rows(idx, owned), owned)
        if diffuse:                                                                       # :337-342
            kept = [torch.as_tensor(entry[1], dtype=torch.int64, device=dev) for entry in patches]
            if kept:
                sum_disjoint(E, group)
                be.sign_strict(E, pts, torch.cat(kept))
        if weights is not None:                                                           # :346-348
            pts[:, 3:] = pts[:, 3:] / weights[:, None]
    return None
