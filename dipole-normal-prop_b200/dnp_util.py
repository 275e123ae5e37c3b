"""Stand-ins for the three helpers of the reference's util.py that the dipole hot
path touches, used ONLY when the reference's own `util` module is not importable
(e.g. on a box without the reference checkout).  When `util` imports, the
drop-in field_utils uses it unchanged, as the reference does.

  gen_grid   util.py:26-36   (the lattice itself is generated in-kernel; this is for callers that want it)
  Transform  util.py:577-609 ('reg' variant: centre by the xyz mean, scale by the largest bounding-box extent)
"""
import torch


def gen_grid(n=10):
    i = torch.arange(0, n ** 3)
    coords = torch.stack([i // (n * n), (i // n) % n, i % n], dim=1).float()
    return (coords / n - 0.5) * 2


class Transform:
    def __init__(self, pc, ttype='reg'):
        if ttype != 'reg':
            raise NotImplementedError("only the 'reg' transform is provided outside the reference's util.py")
        xyz = pc[:, :3]
        self.center = xyz.mean(dim=0)
        self.scale = (xyz.max(dim=0)[0] - xyz.min(dim=0)[0]).max()

    def apply(self, pc):
        out = pc.clone()
        out[:, :3] = (out[:, :3] - self.center[None, :]) / self.scale
        return out

    def inverse(self, pc):
        out = pc.clone()
        out[:, :3] = out[:, :3] * self.scale + self.center[None, :]
        return out

    @staticmethod
    def trans(pc, ttype='reg'):
        t = Transform(pc, ttype=ttype)
        return t.apply(pc), t


# ---------------------------------------------------------------------------- patch partition
# Same names, arguments and nested-list results as util._divide_pc / util.divide_pc / util.merge_nodes
# (util.py:110-150, :338-341, :448-492).  `install()` points the reference's own `util` module at them, so that
# orient_pointcloud.py / orient_large.py / dipole_api.py pick them up through their `util.divide_pc(...)` calls.
def _divide_pc(pc_in: torch.Tensor, n_part: int, ranges=(-1.5, 1.5), min_patch=0):
    """util.py:110-150 on the GPU: one key per point + one radix sort instead of (n_part + 1)^3 mask passes.
    Returns (indices, ijk): indices[c] = [LongTensor of the rows of cell c], ijk[c] = [LongTensor([i, j, k])],
    cells in lexicographic (i, j, k) order, rows ascending -- on pc_in's device, as the reference returns them.
    `min_patch` is unused here by the reference as well (the merge is a separate call)."""
    import dnp_native as _dn
    _dn.require_cuda()
    if pc_in.dtype != torch.float32:
        raise TypeError(f"pc_in must be float32 (got {pc_in.dtype})")
    dev = pc_in.device if pc_in.is_cuda else torch.device("cuda", torch.cuda.current_device())
    perm, offsets, ijk = _dn.voxel_partition(pc_in.to(dev), n_part, ranges)
    sizes = (offsets[1:] - offsets[:-1]).tolist()
    perm, ijk = perm.to(pc_in.device), ijk.to(device=pc_in.device, dtype=torch.int64)
    indices = [[p] for p in torch.split(perm, sizes)] if sizes else []
    return indices, [[c] for c in ijk]


def merge_nodes(pts, indices, ijk, min_patch):
    """util.py:448-492: every patch with fewer than min_patch points is appended to the LAST other patch that owns
    a cell within Chebyshev distance 1 of one of its cells (find_dij keeps overwriting its result), for at most 10
    sweeps; patches still too small are dropped.  Same result as the reference's O(P^2) scan, found through a
    cell -> owner map in O(27 P) per sweep.  Returns (new_indices, new_ijk) like the reference."""
    P = len(ijk)
    segs = [list(indices[i]) for i in range(P)]                     # tensors to concatenate, in merge order
    size = [int(sum(len(t) for t in segs[i])) for i in range(P)]
    cells = [[tuple(int(v) for v in c) for c in ijk[i]] for i in range(P)]
    cell_t = [list(ijk[i]) for i in range(P)]
    owner = {}
    for i in range(P):
        for c in cells[i]:
            owner[c] = i                                             # a cell belongs to exactly one patch
    nbrs = [(a, b, c) for a in (-1, 0, 1) for b in (-1, 0, 1) for c in (-1, 0, 1)]

    remaining, count, max_recursive_merges = True, 0, 10
    while remaining and count < max_recursive_merges:
        remaining = False
        count += 1
        for i in range(P):
            if segs[i] and size[i] < min_patch and cells[i]:
                best = -1
                for (x, y, z) in cells[i]:
                    for (a, b, c) in nbrs:
                        j = owner.get((x + a, y + b, z + c), -1)
                        if j != i and j > best:
                            best = j
                if best != -1:
                    segs[best].extend(segs[i])
                    size[best] += size[i]
                    for c in cells[i]:
                        owner[c] = best
                    cells[best].extend(cells[i])
                    cell_t[best].extend(cell_t[i])
                    segs[i], cells[i], cell_t[i], size[i] = [], [], [], 0
                    if size[best] < min_patch:
                        remaining = True
    if count == max_recursive_merges:
        print('recursive merge failed to merge some patches')

    new_indices, new_ijk = [], []
    for i in range(P):
        if cells[i] and size[i] >= min_patch:
            new_indices.append(torch.cat(segs[i]))
            new_ijk.append(cell_t[i])
    return new_indices, new_ijk


def divide_pc(pc_in: torch.Tensor, n_part: int, ranges=(-1.5, 1.5), min_patch=0):
    """util.py:338-341 as shipped: the cells of _divide_pc, unmerged (the upstream composition that the pipelines
    expect is merge_nodes(pc, *_divide_pc(...), min_patch)[0])."""
    indices, ijk = _divide_pc(pc_in, n_part, ranges, min_patch)
    return indices


def install(util_module=None):
    """Point the reference's `util` module (the one its pipelines import) at the GPU partition:
    util._divide_pc / util.merge_nodes / util.divide_pc.  Returns the module."""
    if util_module is None:
        import util as util_module
    util_module._divide_pc = _divide_pc
    util_module.merge_nodes = merge_nodes
    util_module.divide_pc = divide_pc
    return util_module
