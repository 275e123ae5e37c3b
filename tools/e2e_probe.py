#!/usr/bin/env python
"""Development probe: per-call wall time of the host-buffer entry (the bench's e2e path); set
DNP_HOST_TRACE=1 for the library's own phase timings."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200")); sys.path.insert(0, ROOT)
import torch
import dnp_native as dn
import field_utils
from bench import cloud
P = 1 << 20
src_host, tgt_host = cloud(P, 1, pin=True), cloud(P, 2000, cols=3, pin=True)
for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 12):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    field_utils.field_grad(src_host, tgt_host)
    print("call %d: %.1f ms" % (i, (time.perf_counter() - t0) * 1e3), file=sys.stderr, flush=True)
