#!/usr/bin/env python
"""Constant-bank path probe (development tool): parity against the shared-memory kernel and the float64
oracle, then throughput of both paths.  Each configuration runs in its own process because the path
switches (DNP_CONST_PATH, DNP_CONST_LANES) are read once.  Writes gpurun_out/cpath_probe.jsonl."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
os.makedirs(OUT, exist_ok=True)


def child():
    import numpy as np
    import torch
    sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200"))
    sys.path.insert(0, ROOT)
    import dnp_native as dn

    dev = torch.device("cuda:0")
    tag = os.environ.get("PROBE_TAG", "?")
    log = open(os.path.join(OUT, "cpath_probe.jsonl"), "a")

    def emit(**kw):
        kw["tag"] = tag
        line = json.dumps(kw)
        print(line, flush=True)
        log.write(line + "\n")
        log.flush()

    def cloud(n, seed, cols=6):
        g = torch.Generator().manual_seed(seed)
        xyz = torch.rand(n, 3, generator=g) - 0.5
        if cols == 6:
            nr = torch.randn(n, 3, generator=g)
            xyz = torch.cat([xyz, nr / nr.norm(dim=-1, keepdim=True)], 1)
        return xyz.float().contiguous()

    def timed(fn, reps):
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

    mode = sys.argv[2]
    if mode == "parity":
        S, M = 20000, 140001
        src, tgt = cloud(S, 11), cloud(M, 12, 3)
        tgt[:100] = src[:100, :3]          # coincident pairs
        E = dn.field_grad(src.to(dev), tgt.to(dev), eps=1e-5).cpu().numpy()
        phi = dn.potential(src.to(dev), tgt.to(dev)).cpu().numpy()
        np.save(os.path.join(OUT, f"E_{tag}.npy"), E)
        np.save(os.path.join(OUT, f"phi_{tag}.npy"), phi)
        from oracle import oracle as orc
        sub = np.arange(0, M, 97)[:1500]
        sub[:50] = np.arange(50)
        E64 = orc.field_grad(src.numpy(), tgt.numpy()[sub])
        p64 = orc.potential(src.numpy(), tgt.numpy()[sub])
        eerr = np.abs(E[sub] - E64).max(-1) / (np.abs(E64).max(-1) + 1e-30)
        perr = (np.abs(phi[sub] - p64) / (np.abs(p64) + 1e-5 * np.abs(p64).max()))[50:]
        emit(kind="parity", S=S, M=M, E_rel_max=float(eerr.max()), E_rel_med=float(np.median(eerr)),
             phi_rel_max=float(perr.max()), launches=dn.launch_count())
        # host-buffer entry
        Eh = dn.field_grad_host(src, tgt, eps=1e-5).numpy()
        emit(kind="parity_host", max_abs_diff=float(np.abs(Eh - E).max()), scale=float(np.abs(E).max()))
    elif mode == "speed":
        for S, M in [(1 << 20, 1 << 20), (1 << 18, 1 << 18), (1 << 20, 454656), (1 << 16, 1 << 22)]:
            src, tgt = cloud(S, 1).to(dev), cloud(M, 2, 3).to(dev)
            E = torch.empty(M, 3, device=dev)
            phi = torch.empty(M, device=dev)
            ms = timed(lambda: dn.field_grad(src, tgt, eps=1e-5, out=E), 3)
            emit(kind="field", S=S, M=M, ms=ms, pairs_per_s=S * M / ms * 1e3, frac35=S * M / ms * 1e3 * 35 / 74.45e12)
            ms = timed(lambda: dn.potential(src, tgt, out=phi), 3)
            emit(kind="potential", S=S, M=M, ms=ms, pairs_per_s=S * M / ms * 1e3, frac18=S * M / ms * 1e3 * 18 / 74.45e12)
        srch, tgth = cloud(1 << 20, 1).pin_memory(), cloud(1 << 20, 2, 3).pin_memory()
        import time
        dn.field_grad_host(srch, tgth)
        t0 = time.perf_counter()
        dn.field_grad_host(srch, tgth)
        dt = time.perf_counter() - t0
        emit(kind="field_host_e2e", S=1 << 20, M=1 << 20, ms=dt * 1e3, pairs_per_s=(1 << 40) / dt)


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "child":
        child()
        return
    runs = [("smem", {"DNP_CONST_PATH": "0"}, "parity"),
            ("const", {"DNP_CONST_PATH": "1"}, "parity"),
            ("smem", {"DNP_CONST_PATH": "0"}, "speed"),
            ("const_auto", {"DNP_CONST_PATH": "1"}, "speed"),
            ("const_l1", {"DNP_CONST_PATH": "1", "DNP_CONST_LANES": "1"}, "speed"),
            ("const_l3", {"DNP_CONST_PATH": "1", "DNP_CONST_LANES": "3"}, "speed"),
            ("const_l4", {"DNP_CONST_PATH": "1", "DNP_CONST_LANES": "4"}, "speed")]
    for tag, env, mode in runs:
        e = dict(os.environ)
        e.update(env)
        e["PROBE_TAG"] = tag
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "child", mode], env=e, timeout=900)
        print(f"# {tag} {mode} exit {r.returncode}", flush=True)
    import numpy as np
    try:
        a, b = np.load(os.path.join(OUT, "E_smem.npy")), np.load(os.path.join(OUT, "E_const.npy"))
        pa, pb = np.load(os.path.join(OUT, "phi_smem.npy")), np.load(os.path.join(OUT, "phi_const.npy"))
        line = json.dumps({"kind": "smem_vs_const", "E_max_abs": float(np.abs(a - b).max()),
                           "E_scale": float(np.abs(a).max()),
                           "E_rel_row_max": float((np.abs(a - b).max(-1) / (np.abs(a).max(-1) + 1e-30)).max()),
                           "phi_max_abs": float(np.abs(pa - pb).max()), "phi_scale": float(np.abs(pa).max())})
        print(line)
        open(os.path.join(OUT, "cpath_probe.jsonl"), "a").write(line + "\n")
        for f in ("E_smem.npy", "E_const.npy", "phi_smem.npy", "phi_const.npy"):
            os.remove(os.path.join(OUT, f))
    except OSError as ex:
        print("compare failed", ex)


if __name__ == "__main__":
    main()
