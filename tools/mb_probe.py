#!/usr/bin/env python
"""Pipe-rate / register-file microbenchmarks (development tool): prints lane-ops per SM per clock at 1965 MHz."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200"))
import torch
import dnp_native as dn
info = dn.device_info()
names = {0: "ffma (2 operands from imm/reuse)", 1: "ffma2 (1 reg + 2 uniform)", 2: "mufu.rsq", 3: "field mix",
         4: "ffma2 3 distinct regs", 5: "ffma 3 distinct regs", 6: "ffma2 2 regs + 1 shared", 7: "k7", 8: "k8", 9: "k9"}
for k in [int(a) for a in sys.argv[1:]] or range(7):
    ms, ops = dn.microbench(k, 4000)
    print(json.dumps({"kind": k, "name": names.get(k), "ms": round(ms, 4), "gops": round(ops / ms / 1e6, 1),
                      "per_sm_per_clk_at_1965": round(ops / (ms * 1e-3) / info["sm_count"] / 1.965e9, 2)}), flush=True)
