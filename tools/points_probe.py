"""Per-point greedy propagation (cfg1, tests/golden/points_prop.npz = data/ok.xyz) for each cluster size."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200"))
    import numpy as np, torch, dnp_native as dn
    g = np.load(os.path.join(ROOT, "tests", "golden", "points_prop.npz"))
    n = int(sys.argv[2])
    pc = torch.from_numpy(g["pc_in"][:n].copy()).to("cuda:0")
    best = 1e9
    for _ in range(3):
        w = pc.clone()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); dn.points_propagate(w, diffuse=False); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    print(json.dumps({"ms": best, "us_per_step": best * 1e3 / n}))
    sys.exit(0)
for n in (2000, 10000):
    for c in (0, 1, 2, 4, 8, 16):
        env = dict(os.environ)
        if c: env["DNP_POINTS_CTAS"] = str(c)
        r = subprocess.run([sys.executable, __file__, "child", str(n)], env=env, capture_output=True, text=True)
        print(f"N={n} ctas={c or 'auto'}:", r.stdout.strip().splitlines()[-1] if r.stdout.strip() else r.stderr[-200:], flush=True)
