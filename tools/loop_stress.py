"""Randomised agreement check of the greedy patch loop's schedules (persistent cooperative kernel and slab schedule vs
device-driven launches): random cloud sizes, patch counts (including P > 1024, one-row and empty patches), diffuse on / off, with and
without rest rows.  Prints every disagreement; decisions may only differ where two interactions tie to float32 noise."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200"))
import numpy as np, torch, dnp_native as dn

def run(n_cases=40, verbose=True):
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(11)
    saved = {k: os.environ.get(k) for k in ("DNP_PATCH_LOOP", "DNP_LOOP_MAX_PAIRS")}
    bad = 0
    for case in range(n_cases):
        n = int(rng.integers(40, 60000))
        g = torch.Generator().manual_seed(case)
        d = torch.randn(n, 3, generator=g)
        d = d / d.norm(dim=-1, keepdim=True)
        pc = torch.cat([d * 0.45 + 0.003 * torch.randn(n, 3, generator=g), d], 1).float()
        style = case % 4
        if style == 0:      # spatially coherent patches of similar size
            P = int(rng.integers(1, max(2, n // 50)))
            key = torch.argsort(pc[:, 0] * 7 + pc[:, 1] * 3 + pc[:, 2], stable=True)
            cuts = np.linspace(0, n, P + 1).astype(np.int64)
        elif style == 1:    # ragged: random cut points, duplicates give empty patches
            P = int(rng.integers(2, 400))
            key = torch.randperm(n, generator=g)
            cuts = np.sort(np.concatenate([[0, n], rng.integers(0, n + 1, P - 1)]))
        elif style == 2:    # many tiny patches (P > 1024 when n allows)
            key = torch.randperm(n, generator=g)
            cuts = np.arange(0, n + 1, int(rng.integers(3, 12)))
            if cuts[-1] != n: cuts = np.append(cuts, n)
        else:               # a few big patches (up to the 1024-row limit)
            key = torch.argsort(pc[:, 2], stable=True)
            cuts = np.arange(0, n + 1, int(rng.integers(600, 1024)))
            if cuts[-1] != n: cuts = np.append(cuts, n)
        P = len(cuts) - 1
        if P > 8192 or np.diff(cuts).max() > 1024:
            continue
        flipsel = rng.random(P) < 0.5
        for p in np.nonzero(flipsel)[0]:
            pc[key[cuts[p]:cuts[p + 1]], 3:] *= -1
        diffuse = bool(rng.integers(0, 2))
        use_rest = bool(rng.integers(0, 2))
        offsets = torch.from_numpy(cuts.copy())
        if use_rest:        # representatives = first half of every patch, the rest follows its flips
            rep_cuts = np.concatenate([[0], np.cumsum((np.diff(cuts) + 1) // 2)])
            reps = torch.cat([key[cuts[p]:cuts[p] + (cuts[p + 1] - cuts[p] + 1) // 2] for p in range(P)])
            rest = torch.cat([key[cuts[p] + (cuts[p + 1] - cuts[p] + 1) // 2:cuts[p + 1]] for p in range(P)])
            rest_off = torch.from_numpy(cuts - rep_cuts)
            kw = dict(rest_perm=rest.to(dev), rest_offsets=rest_off)
            perm, offsets = reps.to(dev), torch.from_numpy(rep_cuts)
            kw["tgt_idx"] = perm
        else:
            perm, kw = key.to(dev), {}
        start = int(rng.integers(0, P))
        out = {}
        for mode in ("persistent", "slab", "device"):
            os.environ["DNP_PATCH_LOOP"] = mode
            os.environ["DNP_LOOP_MAX_PAIRS"] = "1e12"
            work = pc.clone().to(dev)
            E = torch.zeros(n, 3, device=dev)
            try:
                order, vals = dn.patch_propagate(work, perm, offsets, start, E, diffuse=diffuse, eps=1e-5, want_trace=True, **kw)
            except dn.DipoleError as e:
                out[mode] = ("error", str(e)[-80:])
                continue
            out[mode] = (order.copy(), vals.copy(), work.cpu(), E.cpu())
        tag = f"case {case}: n={n} P={P} style={style} diffuse={diffuse} rest={use_rest}"
        for mode in ("persistent", "slab"):
            a, b = out[mode], out["device"]
            err_a, err_b = isinstance(a[0], str), isinstance(b[0], str)
            if err_a or err_b:
                if verbose: print(tag, mode, "->", a[:2] if err_a else "", b[:2] if err_b else "")
                bad += err_a
                continue
            same_order = np.array_equal(a[0], b[0])
            same_pts = torch.equal(a[2], b[2])
            rel = ((a[3] - b[3]).norm(dim=-1).max() / b[3].norm(dim=-1).max()).item()
            if mode == "slab" and not diffuse:
                rel = 0.0       # the symmetric form returns no running field (dnp_slab.cu)
            if not (same_order and same_pts) or not rel < 1e-3:
                k = int(np.nonzero(a[0] != b[0])[0][0]) if not same_order else -1
                if verbose: print(tag, mode, f"-> order differs at step {k}: values {a[1][k]:.6e} vs {b[1][k]:.6e}" if k >= 0 else "-> normals or E differ", f"max rel dE {rel:.2e}")
                bad += 1
            else:
                if verbose: print(tag, mode, f"ok (max rel dE {rel:.1e})")
    for k in saved: os.environ.pop(k, None) if saved[k] is None else os.environ.__setitem__(k, saved[k])
    return bad


if __name__ == "__main__":
    print("disagreements:", run(int(sys.argv[1]) if len(sys.argv) > 1 else 40))
