#!/usr/bin/env python
"""Breakdown of the host-buffer (e2e) call: development tool."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200"))
import torch
import dnp_native as dn
P = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
g = torch.Generator().manual_seed(0)
src = torch.rand(P, 6, generator=g).pin_memory()
tgt = torch.rand(P, 3, generator=g).pin_memory()
out_pinned = torch.empty(P, 3).pin_memory()
for name, kw in (("pageable out", {}), ("pinned out", {"out": out_pinned})):
    for i in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        dn.field_grad_host(src, tgt, **kw)
        print(name, i, f"{(time.perf_counter() - t0) * 1e3:.1f} ms", flush=True)
d_src, d_tgt = src.cuda(), tgt.cuda()
E = torch.empty(P, 3, device="cuda")
for i in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    dn.field_grad(d_src, d_tgt, out=E); torch.cuda.synchronize()
    print("device", i, f"{(time.perf_counter() - t0) * 1e3:.1f} ms")
