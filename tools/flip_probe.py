"""Timing probe of the flip solve (dnp_flip.cu) on the B200: exhaustive walk rate and annealing time."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200"))
import dnp_native  # noqa: E402

dev = torch.device("cuda:0")
rng = np.random.default_rng(0)
out = {}


def timed(A, **kw):
    Ad = torch.from_numpy(A).to(dev)
    Bd = -Ad
    for _ in range(2):
        x, obj, st = dnp_native.flip_solve(Ad, Bd, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    x, obj, st = dnp_native.flip_solve(Ad, Bd, **kw)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1), float(obj[0]), float(obj[1]), int(st.item())


for n in (12, 20, 24, 28, 32, 34, 36, 38):
    W = np.triu(rng.normal(size=(n, n)), 1) * (rng.random((n, n)) < 0.4)
    A = W + W.T
    ms, obj, bound, st = timed(A, exact_max_n=40)
    out[f"exhaustive_n{n}"] = {"ms": ms, "states_per_s": 2.0 ** (max(n, 11) - 1) / (ms * 1e-3), "status": st}
    if n <= 36:
        ms2, obj2, _, st2 = timed(A, exact_max_n=0)
        out[f"exhaustive_n{n}"].update(anneal_ms=ms2, anneal_equal=bool(abs(obj2 - obj) <= 1e-9 * abs(obj)))
for n, dens in ((72, 0.1), (300, 0.03), (1000, 0.01), (4096, 0.002)):
    W = np.triu(rng.normal(size=(n, n)), 1) * (rng.random((n, n)) < dens)
    A = W + W.T
    ms, obj, bound, st = timed(A)
    out[f"anneal_n{n}"] = {"ms": ms, "obj_over_bound": obj / bound, "status": st}
print(json.dumps(out, indent=1))
