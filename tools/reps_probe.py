#!/usr/bin/env python
"""Development probe: where the time of the cfg3 stand-in (reps propagation, 1M points, 1000 patches) goes."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200")); sys.path.insert(0, ROOT)
import torch
import dnp_native as dn
import field_utils
from bench import surface_cloud
dev = torch.device("cuda:0")
pc3 = surface_cloud(1 << 20, 10)
key = ((pc3[:, :3] + 0.6) * 8).long().clamp(0, 15)
order = torch.argsort(key[:, 0] * 256 + key[:, 1] * 16 + key[:, 2], stable=True)
chunks = torch.chunk(order, 1000)
reps = [(c[:500].to(dev), c[500:].to(dev)) for c in chunks]
pc3 = pc3.to(dev)
for it in range(3):
    work = pc3.clone()
    torch.cuda.synchronize(); dn.profile_begin(); t0 = time.perf_counter()
    field_utils.strongest_field_propagation_reps(work, reps, diffuse=True)
    torch.cuda.synchronize(); wall = (time.perf_counter() - t0) * 1e3
    k = dn.profile_end(8192)
    print("wall %.1f ms; pairs kernels: %d launches, %.1f ms total, loop-launch median %.3f ms, last %.1f ms" %
          (wall, len(k), sum(k), sorted(k)[len(k) // 2], k[-1]), flush=True)
from torch.profiler import profile, ProfilerActivity
work = pc3.clone()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    field_utils.strongest_field_propagation_reps(work, reps, diffuse=True)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=14, max_name_column_width=60))
print(prof.key_averages().table(sort_by="cpu_time_total", row_limit=14, max_name_column_width=60))
