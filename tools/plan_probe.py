"""Timing of the shared-memory all-pairs kernel under forced plans (DNP_FORCE_TP / DNP_FORCE_NSPLIT), for the
mid-size calls of the greedy drivers (S = a patch, M = all representatives)."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200"))
    import torch, dnp_native as dn
    S, M = int(sys.argv[2]), int(sys.argv[3])
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(1)
    src = torch.randn(S, 6, generator=g).to(dev)
    tgt = torch.randn(M, 6, generator=g).to(dev)
    E = torch.zeros(M, 3, device=dev)
    fn = lambda: dn.field_apply(src, tgt, 1e-5, dn.EPI_ACCUM, out=E)
    for _ in range(3): fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(10): fn()
        b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b) / 10)
    print(json.dumps({"ms": best, "pairs_per_s": S * M / best / 1e-3}))
    sys.exit(0)
SMALL = ((150, 11031), (268, 11031), (300, 300000), (200, 100000), (700, 12000), (1000, 250000), (100, 2000))
if os.environ.get("PLAN_SMALL") == "1":
    for S, M in SMALL:
        for tile, tp, ns in ((0, 0, 0), (256, 1, 1), (256, 2, 1), (64, 0, 1), (64, 0, 2), (64, 0, 3), (64, 0, 5), (64, 0, 8)):
            env = dict(os.environ)
            if tile: env["DNP_FORCE_TILE"] = str(tile)
            if tp: env["DNP_FORCE_TP"] = str(tp)
            if ns: env["DNP_FORCE_NSPLIT"] = str(ns)
            r = subprocess.run([sys.executable, __file__, "child", str(S), str(M)], env=env, capture_output=True, text=True)
            try:
                d = json.loads(r.stdout.strip().splitlines()[-1])
                print(f"S={S} M={M} tile={tile or 'auto'} tp={tp or 'auto'} nsplit={ns or 'auto'}: {d['ms']*1e3:.1f} us  {d['pairs_per_s']:.3e} pairs/s", flush=True)
            except Exception:
                print(f"S={S} M={M} tile={tile} nsplit={ns}: failed {r.stderr[-300:]}", flush=True)
    sys.exit(0)
SHAPES = ((500, 524288), (500, 1 << 20), (200, 100000), (1000, 250000), (150, 11031), (5000, 524288), (2000, 65536),
          (300, 300000), (20000, 200000))
AUTO_ONLY = os.environ.get("PLAN_AUTO_ONLY") == "1"
for S, M in SHAPES:
    for tp in ((0,) if AUTO_ONLY else (0, 4, 2, 1)):
        for ns in ((0,) if tp == 0 else (1, 2, 3, 4)):
            env = dict(os.environ)
            if tp: env["DNP_FORCE_TP"] = str(tp)
            if ns: env["DNP_FORCE_NSPLIT"] = str(ns)
            r = subprocess.run([sys.executable, __file__, "child", str(S), str(M)], env=env, capture_output=True, text=True)
            try:
                d = json.loads(r.stdout.strip().splitlines()[-1])
                print(f"S={S} M={M} tp={tp or 'auto'} nsplit={ns or 'auto'}: {d['ms']:.4f} ms  {d['pairs_per_s']:.3e} pairs/s", flush=True)
            except Exception:
                print(f"S={S} M={M} tp={tp} nsplit={ns}: failed {r.stderr[-200:]}", flush=True)
# the shared-memory kernel at the north-star size (the default path there is the by-value kernel)
if not AUTO_ONLY:
    for tp, ns in ((4, 1), (2, 1), (2, 2), (1, 1)):
        env = dict(os.environ, DNP_FORCE_TP=str(tp), DNP_FORCE_NSPLIT=str(ns))
        r = subprocess.run([sys.executable, __file__, "child", str(1 << 20), str(1 << 20)], env=env, capture_output=True, text=True)
        print(f"S=1M M=1M tp={tp} nsplit={ns}:", r.stdout.strip().splitlines()[-1] if r.stdout.strip() else r.stderr[-200:], flush=True)
