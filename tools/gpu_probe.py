#!/usr/bin/env python
"""First-contact GPU probe: pipe-rate microbenchmarks, parity spot checks against
the oracle, and a sweep of the all-pairs kernels over target-pairs-per-thread.
Writes gpurun_out/probe.jsonl.  Development tool, not part of the product."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200"))
sys.path.insert(0, ROOT)
import dnp_native as dn  # noqa: E402
from oracle import oracle as orc  # noqa: E402

OUT = os.path.join(ROOT, "gpurun_out")
os.makedirs(OUT, exist_ok=True)
log = open(os.path.join(OUT, "probe.jsonl"), "a")


def emit(**kw):
    line = json.dumps(kw)
    print(line, flush=True)
    log.write(line + "\n")
    log.flush()


def cloud(n, seed, dev):
    g = torch.Generator().manual_seed(seed)
    xyz = torch.rand(n, 3, generator=g) - 0.5
    nr = torch.randn(n, 3, generator=g)
    nr = nr / nr.norm(dim=-1, keepdim=True)
    return torch.cat([xyz, nr], 1).float().to(dev)


def timeit(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def main():
    dev = torch.device("cuda:0")
    info = dn.device_info()
    emit(kind="device", name=torch.cuda.get_device_name(0), **info)
    what = sys.argv[1:] or ["micro", "parity", "sweep", "points"]

    if "micro" in what:
        names = {0: "ffma", 1: "ffma2", 2: "mufu_rsq", 3: "field_mix"}
        for k, nm in names.items():
            ms, ops = dn.microbench(k, 4000)
            emit(kind="microbench", name=nm, ms=ms, gops=ops / ms / 1e6,
                 per_sm_per_clk_at_1965=ops / (ms * 1e-3) / info["sm_count"] / 1.965e9)

    if "parity" in what:
        for (S, M) in [(1, 1), (7, 5), (255, 257), (1000, 333), (5000, 3000)]:
            src, tgt = cloud(S, 100 + S, dev), cloud(M, 200 + M, dev)
            for tp in (1, 2, 4):
                os.environ["DNP_FORCE_TP"] = str(tp)
                for ns in (1, 3):
                    os.environ["DNP_FORCE_NSPLIT"] = str(ns)
                    E = dn.field_grad(src, tgt).cpu().numpy()
                    ph = dn.potential(src, tgt).cpu().numpy()
                    E64 = orc.field_grad(src.cpu().numpy(), tgt.cpu().numpy())
                    p64 = orc.potential(src.cpu().numpy(), tgt.cpu().numpy())
                    eE = np.abs(E - E64).max() / max(np.abs(E64).max(), 1e-30)
                    eP = np.abs(ph - p64).max() / max(np.abs(p64).max(), 1e-30)
                    emit(kind="parity", S=S, M=M, tp=tp, nsplit=ns, E_relmax=float(eE), phi_relmax=float(eP))
        os.environ.pop("DNP_FORCE_TP", None)
        os.environ.pop("DNP_FORCE_NSPLIT", None)
        # coincident pairs
        cl = cloud(300, 7, dev)
        E = dn.field_grad(cl, cl).cpu().numpy()
        E64 = orc.field_grad(cl.cpu().numpy(), cl.cpu().numpy())
        emit(kind="parity_self", E_relmax=float(np.abs(E - E64).max() / np.abs(E64).max()),
             finite=bool(np.isfinite(E).all()))
        ph = dn.potential(cl, cl).cpu().numpy()
        p64 = orc.potential(cl.cpu().numpy(), cl.cpu().numpy(), drop_coincident=True)
        emit(kind="parity_self_phi", relmax=float(np.abs(ph - p64).max() / np.abs(p64).max()))
        m = dn.lattice_mean_potential(cl).item()
        emit(kind="lattice", gpu=m, oracle=orc.measure_mean_potential(cl.cpu().numpy()))

    if "sweep" in what:
        for n in (65536, 262144, 1048576):
            src, tgt = cloud(n, 1, dev), cloud(n, 2, dev)
            E = torch.empty(n, 3, device=dev)
            phi = torch.empty(n, device=dev)
            for tp in (1, 2, 4):
                os.environ["DNP_FORCE_TP"] = str(tp)
                ms = timeit(lambda: dn.field_grad(src, tgt, out=E), reps=3 if n < 1000000 else 2)
                emit(kind="sweep", op="field", n=n, tp=tp, ms=ms, pairs_per_s=n * n / ms * 1e3)
                ms = timeit(lambda: dn.potential(src, tgt, out=phi), reps=3 if n < 1000000 else 2)
                emit(kind="sweep", op="potential", n=n, tp=tp, ms=ms, pairs_per_s=n * n / ms * 1e3)
            os.environ.pop("DNP_FORCE_TP", None)
            ms = timeit(lambda: dn.field_grad(src, tgt, out=E), reps=2)
            emit(kind="sweep", op="field", n=n, tp="auto", ms=ms, pairs_per_s=n * n / ms * 1e3)

    if "points" in what:
        g = np.load(os.path.join(ROOT, "tests", "golden", "points_prop.npz"))
        pc = torch.from_numpy(g["pc_in"])
        for tag in ("n512", "n2000", "full"):
            n = int(g[f"{tag}_n"])
            pts = pc[:n].clone().to(dev)
            t0 = time.time()
            order, fl = dn.points_propagate(pts, diffuse=bool(g[f"{tag}_diffuse"]), want_trace=True)
            torch.cuda.synchronize()
            dt = time.time() - t0
            emit(kind="points", tag=tag, n=n, seconds=dt,
                 order_eq=float((order.cpu().numpy() == g[f"{tag}_order"]).mean()),
                 flips_eq=float((fl.cpu().numpy().astype(bool) == g[f"{tag}_flipped"]).mean()),
                 normals_eq=bool((pts[:, 3:].cpu().numpy() == g[f"{tag}_normals"]).all()))
            pts = pc[:n].clone().to(dev)
            ms = timeit(lambda: dn.points_propagate(pts), reps=2)
            emit(kind="points_time", tag=tag, n=n, ms=ms, us_per_step=ms * 1e3 / n)


if __name__ == "__main__":
    main()
