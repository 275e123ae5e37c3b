"""Lattice sign test timing (development tool): the mean-only kernel of dnp_lattice.cu against the generic
all-pairs path (DNP_LATTICE_FAST=0), CUDA events, best of 5 x 20 calls."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200"))
import torch  # noqa: E402
import dnp_native as dn  # noqa: E402

dev = torch.device("cuda:0")
for S in (10000, 100000, 1 << 20, 1 << 22):
    g = torch.Generator().manual_seed(3)
    pc = torch.cat([torch.rand(S, 3, generator=g) - 0.5, torch.randn(S, 3, generator=g)], 1).to(dev)
    row = {"S": S}
    for name, env in (("fast", "1"), ("generic", "0")):
        os.environ["DNP_LATTICE_FAST"] = env
        for _ in range(3):
            m = dn.lattice_mean_potential(pc)
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(5):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(20):
                m = dn.lattice_mean_potential(pc)
            b.record()
            torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b) / 20)
        row[name] = {"ms": best, "pairs_per_s": S * 1000 / best * 1e3, "mean": m.item()}
    print(json.dumps(row), flush=True)
