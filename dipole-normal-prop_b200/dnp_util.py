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


def merge_cells(size0, cells, n_part, min_patch):
    """The sequential merge rule of util.merge_nodes (util.py:448-492) over integer cell ids: size0[c] points in
    cell c, cells[c] = (i, j, k).  Returns (kept patch ids, members[patch] = its cells in merge order, sizes)."""
    P0 = len(size0)
    side = int(n_part) + 4                                           # cell coordinates shifted by 1: neighbours stay >= 0
    key_of = [((int(c[0]) + 1) * side + (int(c[1]) + 1)) * side + (int(c[2]) + 1) for c in cells]
    owner = dict(zip(key_of, range(P0)))                             # cell key -> patch currently owning it
    deltas = [(a * side + b) * side + c for a in (-1, 0, 1) for b in (-1, 0, 1) for c in (-1, 0, 1)]
    members = [[i] for i in range(P0)]                               # cells of patch i, in merge order
    size = list(size0)
    remaining, sweeps = True, 0
    while remaining and sweeps < 10:                                 # util.py:463-466: at most 10 sweeps
        remaining = False
        sweeps += 1
        for i in range(P0):
            if members[i] and size[i] < min_patch:
                best = -1
                for c in members[i]:
                    k = key_of[c]
                    for d in deltas:
                        j = owner.get(k + d, -1)
                        if j != i and j > best:                      # find_dij keeps the LAST neighbouring patch
                            best = j
                if best != -1:
                    members[best].extend(members[i])
                    size[best] += size[i]
                    for c in members[i]:
                        owner[key_of[c]] = best
                    members[i], size[i] = [], 0
                    if size[best] < min_patch:
                        remaining = True
    if sweeps == 10:
        print('recursive merge failed to merge some patches')
    kept = [i for i in range(P0) if members[i] and size[i] >= min_patch]
    return kept, members, size


def partition_patches(pc_in: torch.Tensor, n_part: int, ranges=(-1.5, 1.5), min_patch=0, want_ijk=False):
    """merge_nodes(pc, *_divide_pc(pc[:, :3], n_part), min_patch) -- the patches the pipelines work on -- without
    materialising one tensor per CELL: the voxel partition stays a CSR on the device, the merge (same sequential
    rule, util.py:448-492) runs on the host over integer cell ids, and the merged patches are gathered with one
    device pass.  At --number_parts 41 a 1M-point cloud has several thousand occupied cells but ~1000 patches.
    Returns the list of per-patch row tensors (and with want_ijk the per-patch lists of cells [i, j, k])."""
    import numpy as np
    import dnp_native as _dn
    _dn.require_cuda()
    if pc_in.dtype != torch.float32:
        raise TypeError(f"pc_in must be float32 (got {pc_in.dtype})")
    dev = pc_in.device if pc_in.is_cuda else torch.device("cuda", torch.cuda.current_device())
    perm, offsets, ijk = _dn.voxel_partition(pc_in[:, :3].to(dev), n_part, ranges)
    off = offsets.cpu().numpy()
    cells = ijk.cpu().numpy().astype(np.int64)
    P0 = cells.shape[0]
    if P0 == 0:
        return ([], []) if want_ijk else []
    size0 = (off[1:] - off[:-1]).tolist()
    kept, members, size = merge_cells(size0, cells, n_part, min_patch)
    if not kept:
        return ([], []) if want_ijk else []
    order = np.fromiter((c for i in kept for c in members[i]), dtype=np.int64)
    counts = torch.from_numpy(np.asarray(size0, dtype=np.int64)[order]).to(dev)
    starts = offsets[:-1][torch.from_numpy(order).to(dev)]
    pos = torch.cumsum(counts, 0) - counts
    total = int(sum(size[i] for i in kept))
    rep = torch.repeat_interleave(torch.arange(order.shape[0], device=dev), counts, output_size=total)
    rows = perm[torch.arange(total, device=dev) - pos[rep] + starts[rep]].to(pc_in.device)
    patches = list(torch.split(rows, [size[i] for i in kept]))
    if not want_ijk:
        return patches
    cells_t = torch.from_numpy(cells)
    return patches, [[cells_t[c] for c in members[i]] for i in kept]


def divide_pc(pc_in: torch.Tensor, n_part: int, ranges=(-1.5, 1.5), min_patch=0):
    """util.divide_pc as the pipelines USE it (orient_pointcloud.py:25-27, orient_large.py:33-35, dipole_api.py:38-40:
    a list of per-patch row tensors, cloned and indexed): the upstream composition merge_nodes(pc, *_divide_pc(...),
    min_patch)[0].  The fork's util.py:338-341 returns the unmerged nested cell lists instead, on which those very
    call sites raise -- there is no working behaviour of the shipped function to keep."""
    return partition_patches(pc_in, n_part, ranges, min_patch)


def install(util_module=None):
    """Point the reference's `util` module (the one its pipelines import) at the GPU path: the voxel partition
    (_divide_pc / merge_nodes / divide_pc), the patch graph (_lzd_divide_pc / lzd_merge_nodes / divide_pc_to_graph)
    and the text IO (xyz2tensor / export_pc).  Returns the module."""
    if util_module is None:
        import util as util_module
    for name in ("_divide_pc", "merge_nodes", "divide_pc", "_lzd_divide_pc", "lzd_merge_nodes", "divide_pc_to_graph",
                 "xyz2tensor", "export_pc"):
        setattr(util_module, name, globals()[name])
    return util_module


# ---------------------------------------------------------------------------- patch graph
# Host bookkeeping around the flip-score kernels, same names and results as the reference's util.py.  The
# reference's timers, normal estimation (open3d / torch_cluster) and orient_center are NOT restated here: its
# pipelines keep using their own util module (dnp_dropin.activate), which is where those live.
def npxyz2tensor(np_pc, append_normals=True):
    """util.py:71-77: numpy [N, 3|6] -> tensor of the same dtype (float64 from the wire), zero normals appended."""
    import numpy as np
    if np_pc.shape[1] == 3 and append_normals:
        np_pc = np.concatenate([np_pc, np.zeros((np_pc.shape[0], 3))], axis=1)
    return torch.tensor(np_pc)


def knn_pca_normals(xyz, k=30):
    """Unoriented normals for clouds that arrive without any (the server's wire format, socket_server_para.py:145-166)
    when the reference's util.estimate_normals (open3d, util.py:551-567) is unavailable: direction of least variance
    of each point's k nearest neighbours.  Plumbing in front of the dipole path (sign arbitrary, the propagation
    orients it); second moments are accumulated in float64 on the device."""
    pts = xyz[:, :3]
    n = pts.shape[0]
    k = max(1, min(int(k), n - 1))
    p64 = pts.double()
    normals = torch.empty_like(pts)
    rows = max(1, min(n, (1 << 26) // max(n, 1)))
    for lo in range(0, n, rows):
        block = p64[lo:lo + rows]
        dist = torch.cdist(block, p64)
        dist[torch.arange(block.shape[0], device=pts.device), torch.arange(lo, lo + block.shape[0], device=pts.device)] = float("inf")
        nb = p64[dist.topk(k, dim=1, largest=False).indices]               # [rows, k, 3], self excluded
        mean = nb.mean(dim=1)
        second = torch.einsum("rki,rkj->rij", nb, nb) / k - mean[:, :, None] * mean[:, None, :]
        normals[lo:lo + rows] = torch.linalg.eigh(second).eigenvectors[:, :, 0].to(pts.dtype)
    return torch.cat([pts, normals], dim=1)


def _lzd_divide_pc(pc_in: torch.Tensor, n_part: int, ranges=(-1.5, 1.5), min_patch=0):
    """util.py:85-108: cell of every point by searchsorted on linspace(lo, hi, n_part + 1) (left-closed cells,
    key = x + y n + z n^2), patches in ascending key order with their rows ascending.  The reference groups the
    rows with a Python loop over all points (:100-101); here one stable sort by key.  Returns (indices, patch_ijk)
    with the reference's nesting: indices[c] = [LongTensor], patch_ijk[c] = [LongTensor([x, y, z])] (the cell of
    the patch's first point, :106)."""
    dev = pc_in.device
    edges = torch.linspace(ranges[0], ranges[1], n_part + 1).to(device=dev, dtype=pc_in.dtype)
    cell = torch.stack([torch.searchsorted(edges, pc_in[:, a].contiguous(), right=True) - 1 for a in range(3)], dim=1)
    key = cell[:, 0] + cell[:, 1] * n_part + cell[:, 2] * n_part * n_part
    skey, rows = torch.sort(key, stable=True)
    _, counts = torch.unique_consecutive(skey, return_counts=True)
    parts = torch.split(rows, counts.tolist())
    indices = [[p] for p in parts]
    patch_ijk = [[cell[p[0]]] for p in parts]
    return indices, patch_ijk


def _manhattan_neighbours(c):
    x, y, z = c
    return ((x - 1, y, z), (x + 1, y, z), (x, y - 1, z), (x, y + 1, z), (x, y, z - 1), (x, y, z + 1))


def lzd_merge_nodes(pts, indices, ijk, min_patch):
    """util.py:418-447: one pass over the cells in order; a cell with at most min_patch points is appended to the
    neighbouring patch (a source cell at Manhattan distance 1) that currently counts the fewest points -- counts
    as the reference keeps them, including its bookkeeping quirk of crediting the moved points to the FIRST
    neighbour (:435).  The reference's O(P^2 |cells|^2) neighbour scan is a cell -> owner map here.
    Returns (new_indices, new_ijk, ijk_source)."""
    import numpy as np
    ijk = [c[0] for c in ijk]
    indices = [c[0] for c in indices]
    n0 = len(indices)
    cell_of = [tuple(int(v) for v in c) for c in ijk]
    by_cell = {}
    for i, c in enumerate(cell_of):
        by_cell.setdefault(c, []).append(i)
    owner = list(range(n0))                      # current patch of every original cell
    sources = [[i] for i in range(n0)]
    pt_count = np.array([len(i) for i in indices])
    for i in range(n0):
        count = pt_count[i]
        if count > min_patch:
            continue
        near = set()
        for s in sources[i]:
            for c in _manhattan_neighbours(cell_of[s]):
                for o in by_cell.get(c, ()):
                    if owner[o] != i:
                        near.add(owner[o])
        if not near:
            continue
        neighbor = np.array(sorted(near))
        smallest = int(neighbor[np.argmin(pt_count[neighbor])])
        for s in sources[i]:
            owner[s] = smallest
        sources[smallest] += sources[i]
        sources[i] = []
        pt_count[neighbor[0]] += count
        pt_count[i] = 0
    new_indices, new_ijk, ijk_source = [], [], []
    for i in range(n0):
        if not sources[i]:
            continue
        new_indices.append(torch.cat([indices[j] for j in sources[i]]))
        new_ijk.append(ijk[i])
        ijk_source.append([ijk[j] for j in sources[i]])
    return new_indices, new_ijk, ijk_source


def graph_edges(ijk_source):
    """Patch pairs (i < j, lexicographic) owning two cells at Manhattan distance 1: the pairs util.py:199-203
    keeps, found through a cell map instead of the all-pairs scan."""
    cell_patch = {}
    for p, cells in enumerate(ijk_source):
        for c in cells:
            cell_patch.setdefault(tuple(int(v) for v in c), set()).add(p)
    edges = set()
    for p, cells in enumerate(ijk_source):
        for c in cells:
            for nb in _manhattan_neighbours(tuple(int(v) for v in c)):
                for q in cell_patch.get(nb, ()):
                    if q > p:
                        edges.add((p, q))
    return sorted(edges)


def divide_pc_to_graph(pc_in: torch.Tensor, n_part: int, ranges=(-1.5, 1.5), min_patch=0, edge_calculator=None,
                       point_estimator=None):
    """util.py:160-212.  Grid patches -> per-patch orientation (point_estimator, in place on pc_in) -> BidGraph
    with one flip score per adjacent patch pair.  When edge_calculator is this package's
    field_utils.field_edge_calculator, all scores come from ONE launch of the batched edge kernel
    (dnp_edge_interactions_batched_f32) instead of one call per edge.  Returns (G, indices)."""
    import numpy as np
    import field_utils
    from graph import BidGraph, BiEdge
    indices, ijk = _lzd_divide_pc(pc_in, n_part, ranges, min_patch)
    indices, ijk, ijk_source = lzd_merge_nodes(pc_in, indices, ijk, min_patch)
    if point_estimator is not None:
        import dnp_native as _dn
        batched = getattr(point_estimator, "batched", None)
        if (batched is not None and pc_in.is_cuda and pc_in.dtype == torch.float32 and indices
                and max(len(i) for i in indices) <= _dn.load().dnp_points_batched_max_n()):
            batched(pc_in, indices)  # every patch in one launch (graph_dipole.st_propagation_points_batched)
        else:
            for idx in indices:      # the reference starts one Python thread per patch; the device serialises them
                pc_in[idx] = point_estimator(pc_in[idx])
    G = BidGraph()
    G.V = [i for i in range(len(indices))]
    pairs = graph_edges(ijk_source)
    assert edge_calculator is not None or not pairs
    if pairs and edge_calculator is field_utils.field_edge_calculator:
        w, invw = field_utils.field_edge_calculator_batched(pc_in, indices, np.asarray(pairs, dtype=np.int64))
        for (i, j), a, b in zip(pairs, w, invw):
            G.E.append(BiEdge(i, j, np.asarray(a), b))
    else:
        for i, j in pairs:
            w, invw = edge_calculator(pc_in[indices[i]], pc_in[indices[j]])
            G.E.append(BiEdge(i, j, w, invw))
    return G, indices


# ---------------------------------------------------------------------------- text IO + wire format
def xyz2tensor_cuda(txt, append_normals=True, device=None):
    """util.py:53-69 with the result left on the GPU.  `txt`: str or bytes.  Line splitting, the "nan" filter, the
    single-space tokenisation, float() rounding and the float32 cast all happen in dnp_xyz_parse_f32."""
    import dnp_native as _dn
    raw = txt.encode("utf-8", errors="surrogateescape") if isinstance(txt, str) else txt
    rows, mask, bad, _ = _dn.xyz_parse(raw, device=device)
    if bad >= 0:
        line = bytes(raw).split(b"\n")[bad].strip()
        raise ValueError(f"could not convert string to float in line {bad}: {line[:80]!r}")
    if rows.shape[0] == 0:
        raise RuntimeError("stack expects a non-empty TensorList")             # torch.stack([]) in the reference
    if not append_normals:
        if mask == 3:
            raise RuntimeError("stack expects each tensor to be equal size (3- and 6-column lines mixed)")
        if mask == 1:
            return rows[:, :3].contiguous()
    return rows


def xyz2tensor(txt, append_normals=True):
    """util.py:53-69: the parsed cloud as a CPU float32 tensor [n, 6] (or [n, 3]), like the reference returns it."""
    return xyz2tensor_cuda(txt, append_normals).cpu()


def export_text(pc, as_view=False):
    """The text util.export_pc writes for pc [C, N] (it iterates pc.transpose(0, 1)): bytes, no trailing newline
    (as_view: a memoryview of the downloaded buffer instead, see dnp_native.xyz_format)."""
    import dnp_native as _dn
    _dn.require_cuda()
    if pc.dtype != torch.float32:
        raise TypeError(f"export_pc formats float32 clouds (got {pc.dtype})")
    dev = pc.device if pc.is_cuda else torch.device("cuda", torch.cuda.current_device())
    return _dn.xyz_format(pc.detach().to(dev).transpose(0, 1), as_view=as_view)


def export_pc(pc, dest):
    """util.py:46-51.  The text is formatted on the device and streamed into the file by the library
    (dnp_xyz_export_fd_f32: pinned staging, the download of a chunk overlaps the write of the previous one)."""
    import dnp_native as _dn
    _dn.require_cuda()
    if pc.dtype != torch.float32:
        raise TypeError(f"export_pc formats float32 clouds (got {pc.dtype})")
    dev = pc.device if pc.is_cuda else torch.device("cuda", torch.cuda.current_device())
    with open(dest, 'wb+') as file:
        _dn.xyz_export_fd(pc.detach().to(dev).transpose(0, 1), file.fileno())


def wire_decode(payload, device=None):
    """socket_server_para.py:145-166: the request's N x 3 float64 payload -> [N, 6] float32 rows on the GPU."""
    import dnp_native as _dn
    return _dn.wire_decode(payload, device=device)


def wire_encode(pc):
    """socket_server_para.py:191: result.astype(np.float64).tobytes() for a [N, 6] float32 cloud."""
    import dnp_native as _dn
    dev = pc.device if pc.is_cuda else torch.device("cuda", torch.cuda.current_device())
    return _dn.wire_encode(pc.to(dev))
