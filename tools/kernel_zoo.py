"""One small workload per kernel of the hot path, for profiling (`ncu -k regex:<kernel> ... python tools/kernel_zoo.py <name>`)
and for a plain timing table (`python tools/kernel_zoo.py all` prints CUDA-event times, best of 3).  Development tool."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200"))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import dnp_native as dn  # noqa: E402
import field_utils  # noqa: E402
from bench import cloud, surface_cloud  # noqa: E402

dev = torch.device("cuda:0")


def patches_1000(pc):
    key = ((pc[:, :3] + 0.6) * 8).long().clamp(0, 15)
    order = torch.argsort(key[:, 0] * 256 + key[:, 1] * 16 + key[:, 2], stable=True)
    chunks = torch.chunk(order, 1000)
    offsets = torch.zeros(1001, dtype=torch.int64)
    offsets[1:] = torch.tensor([c.numel() for c in chunks]).cumsum(0)
    return torch.cat(chunks).to(dev), offsets.to(dev), chunks


def make(name):
    """returns (callable, work description, unit count, unit name)"""
    n = 1 << 20
    if name in ("field", "potential"):
        src, tgt = cloud(n, 1).to(dev), cloud(n, 2, cols=3).to(dev)
        out = torch.empty(n, 3 if name == "field" else 1, device=dev)
        if name == "field":
            return (lambda: dn.field_grad(src, tgt, out=out)), "field_grad 1M x 1M (upairs_tile_kernel<false>)", float(n) * n, "pairs"
        return (lambda: dn.potential(src, tgt, out=out.view(-1))), "potential 1M x 1M (upairs_tile_kernel<true>)", float(n) * n, "pairs"
    if name in ("shares", "fused"):
        S = 5000 if name == "shares" else 500
        src, tgt = cloud(S, 1).to(dev), cloud(524288, 2, cols=3).to(dev)
        out = torch.empty(524288, 3, device=dev)
        return (lambda: dn.field_grad(src, tgt, out=out)), f"field_grad {S} x 524288 (upairs_kernel<false,4>)", S * 524288.0, "pairs"
    if name == "smem":
        src, tgt = cloud(268, 1).to(dev), cloud(11031, 2, cols=3).to(dev)
        out = torch.empty(11031, 3, device=dev)
        return (lambda: dn.field_grad(src, tgt, out=out)), "field_grad 268 x 11031 (pairs_kernel, fandisk step)", 268 * 11031.0, "pairs"
    if name == "lattice":
        src = cloud(n, 1).to(dev)
        return (lambda: dn.lattice_mean_potential(src)), "lattice mean, 1M sources x 1000 nodes (lattice_sum_kernel)", n * 1000.0, "pairs"
    if name in ("points", "points_grid"):
        N = 10000 if name == "points" else 120000
        pc = surface_cloud(N, 7).to(dev)
        return (lambda: field_utils.strongest_field_propagation_points(pc.clone())), f"point propagation N={N}", float(N), "steps"
    if name in ("edges", "patch_inter"):
        pc = surface_cloud(n, 10)
        perm, offsets, chunks = patches_1000(pc)
        pcd = pc.to(dev)
        if name == "edges":
            e = [(i, (i + d) % 1000) for i in range(1000) for d in (1, 2, 3)]
            edges = torch.tensor(e + [(j, i) for i, j in e]).to(dev)
            pairs = float(sum(chunks[i].numel() * chunks[j].numel() for i, j in edges.tolist()))
            return (lambda: dn.edge_interactions_batched(pcd, perm, offsets, edges)), "6000 patch-pair edges of ~1048 x 1048 (edge_interactions_kernel)", pairs, "pairs"
        E = torch.randn(n, 3, device=dev)
        return (lambda: dn.patch_interactions(E, pcd, perm, offsets)), "per-patch sum E.n, 1M rows / 1000 patches (patch_interactions_kernel)", n * 36.0, "bytes"
    if name == "sign":
        big = torch.randn(1 << 24, 6, device=dev)
        Eb = torch.randn(1 << 24, 3, device=dev)
        return (lambda: dn.field_sign(Eb, big)), "diffuse sign of 16M rows (field_sign_kernel)", (1 << 24) * 36.0, "bytes"
    if name == "xie":
        pc = surface_cloud(10000, 11).to(dev)
        g = torch.Generator().manual_seed(12)
        orders = torch.stack([torch.randperm(10000, generator=g) for _ in range(5)]).to(dev)
        return (lambda: field_utils.xie_propagation_points_in_order(pc, 1e-2, orders, diffuse=True)), "xie in-order, 10000 pts x 5 chains (xie_chain_kernel)", 50000.0, "steps"
    if name == "partition":
        pc = surface_cloud(n, 10).to(dev)
        return (lambda: dn.voxel_partition(pc, 41)), "voxel partition 1M pts, 41 parts (cell_key_kernel + CUB)", n * 20.0, "bytes"
    if name == "slab":
        pc = surface_cloud(n, 10)
        _, _, chunks = patches_1000(pc)
        reps = [c[:500].to(dev) for c in chunks]
        perm = torch.cat(reps)
        offsets = torch.arange(1001, dtype=torch.int64) * 500
        rest = torch.cat([c[500:] for c in chunks]).to(dev)
        rest_off = torch.tensor([0] + [c.numel() - 500 for c in chunks], dtype=torch.int64).cumsum(0)
        pcd = pc.to(dev)

        def run():
            work, E = pcd.clone(), torch.zeros(n, 3, device=dev)
            dn.patch_propagate(work, perm, offsets, 0, E, diffuse=True, rest_perm=rest, rest_offsets=rest_off, tgt_idx=perm)
        return run, "greedy loop, 1000 patches x 500 representatives, diffuse (slab schedule: upairs_tile_kernel<false,4,true> + slab_*)", 500.0 * 500000 * 1000, "pairs"
    raise SystemExit(f"unknown workload {name}")


NAMES = ["field", "potential", "shares", "fused", "smem", "lattice", "points", "points_grid", "edges", "patch_inter", "sign",
         "xie", "partition", "slab"]

if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    for name in (NAMES if which == "all" else [which]):
        fn, what, units, unit = make(name)
        fn()
        torch.cuda.synchronize()
        best = 1e30
        for _ in range(3 if which == "all" else 2):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b))
        print(json.dumps({"workload": name, "what": what, "ms": best, unit + "_per_s": units / best * 1e3}), flush=True)
