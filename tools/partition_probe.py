#!/usr/bin/env python
"""Development probe: voxel partition throughput and merge time."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200")); sys.path.insert(0, ROOT)
import torch
import dnp_native as dn, dnp_util
from bench import surface_cloud
dev = torch.device("cuda:0")
lib = dn.load()
for n in (1 << 20, 1 << 24):
    pc = surface_cloud(n, 5).to(dev)
    N = pc.shape[0]
    perm = torch.empty(N, dtype=torch.int64, device=dev); offsets = torch.zeros(N + 1, dtype=torch.int64, device=dev)
    ijk = torch.empty(N, 3, dtype=torch.int32, device=dev); counts = torch.zeros(2, dtype=torch.int64, device=dev)
    nb = lib.dnp_voxel_partition_workspace_bytes(N); ws = torch.empty(nb, dtype=torch.uint8, device=dev)
    def call():
        dn._check(lib.dnp_voxel_partition_f32(dn._ptr(pc), N, 6, 41, -1.5, 1.5, dn._ptr(perm), dn._ptr(offsets), dn._ptr(ijk),
                                              dn._ptr(counts), dn._ptr(ws), nb, dn._stream(pc)))
    call(); torch.cuda.synchronize(); best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True); e0.record(); call(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print(f"N={n}: ABI call {best:.3f} ms = {n / best / 1e6:.2f} G points/s, {counts.tolist()} (cells, in-box)", flush=True)
    t0 = time.perf_counter(); idx, cells = dnp_util._divide_pc(pc[:, :3], 41); torch.cuda.synchronize(); t1 = time.perf_counter()
    merged, _ = dnp_util.merge_nodes(pc, idx, cells, 100); t2 = time.perf_counter()
    print(f"   python _divide_pc {1e3 * (t1 - t0):.1f} ms ({len(idx)} cells), merge_nodes {1e3 * (t2 - t1):.1f} ms -> {len(merged)} patches", flush=True)
