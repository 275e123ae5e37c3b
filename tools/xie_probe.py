#!/usr/bin/env python
"""Development probe: timings of the xie kernels (matrix, in-order propagation, diffuse mat-vec)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200")); sys.path.insert(0, ROOT)
import torch
import dnp_native as dn
from bench import surface_cloud
dev = torch.device("cuda:0")
def best(fn, reps=3):
    fn(); torch.cuda.synchronize(); b = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); b = min(b, e0.elapsed_time(e1))
    return b
for n, T in [(1500, 5), (10000, 5), (10000, 15), (30000, 5), (100000, 5)]:
    pc = surface_cloud(n, 3).to(dev)
    g = torch.Generator().manual_seed(1)
    orders = torch.stack([torch.randperm(n, generator=g) for _ in range(T)]).to(dev)
    ms0 = best(lambda: dn.xie_propagate_in_order(pc, orders, diffuse=False), 2)
    ms1 = best(lambda: dn.xie_propagate_in_order(pc, orders, diffuse=True), 2)
    line = f"N={n} T={T}: chains {ms0:.2f} ms ({ms0 / n * 1e3:.2f} us/step), with diffuse {ms1:.2f} ms"
    if n <= 30000:
        ms2 = best(lambda: dn.xie_matrix(pc, pc), 2)
        line += f"; matrix {ms2:.2f} ms = {n * n * 4 / ms2 / 1e6:.0f} GB/s written"
    print(line, flush=True)
pc = surface_cloud(20000, 3).to(dev)
print("knn_radius2 20000 k=40: %.2f ms" % best(lambda: dn.knn_radius2(pc, pc, 40), 2))
