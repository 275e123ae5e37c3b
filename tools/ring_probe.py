"""A/B timing of the all-pairs paths on one GPU (development tool): the constant-window kernels (dnp_upairs.cu,
DNP_RING and its DNP_RING_* knobs) and the shared-memory kernel (dnp_pairs.cu, DNP_RING=0).  CUDA events around the native call, best of 5 groups.  Prints one JSON line per case."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200"))
import torch  # noqa: E402
import dnp_native as dn  # noqa: E402

dev = torch.device("cuda:0")


def setenv(**kw):
    for k in list(os.environ):
        if k.startswith("DNP_"):
            os.environ.pop(k, None)
    for k, v in kw.items():
        os.environ[k] = str(v)


def time_call(fn, reps):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b) / reps)
    return best


EXTRA = [a for a in sys.argv[2:] if "=" in a]
ONLY = set(os.environ.get("PROBE_ONLY", "").split(",")) - {""}
KEYS = set()


def case(S, M, potential=False):
    g = torch.Generator().manual_seed(1)
    src = torch.cat([torch.rand(S, 3, generator=g) - 0.5, torch.randn(S, 3, generator=g)], 1).to(dev)
    tgt = (torch.rand(M, 3, generator=g) - 0.5).to(dev)
    out = torch.zeros(M, 1 if potential else 3, device=dev)
    if potential:
        fn = lambda: dn.potential(src, tgt, out=out.view(-1))
    else:
        fn = lambda: dn.field_apply(src, tgt, 1e-5, dn.EPI_STORE, out=out)
    pairs = S * M
    reps = max(1, min(20, int(2e11 / pairs)))
    res = {}
    configs = [("window", dict(DNP_RING=2)), ("window_tp1", dict(DNP_RING=2, DNP_RING_TP=1)),
               ("smem", dict(DNP_RING=0))]
    for spec in EXTRA:                      # name=ENV:VAL,ENV:VAL from the command line
        name, envs = spec.split("=", 1)
        configs.append((name, dict(e.split(":") for e in envs.split(","))))
    if ONLY:
        configs = [c for c in configs if c[0] in ONLY]
    ref = None
    for name, env in configs:
        if name == "window_tp1" and M > 200000:
            continue
        setenv(**env)
        try:
            ms = time_call(fn, reps)
            val = out.clone()
            if ref is None:
                ref = val
            dmax = float((val - ref).abs().max() / ref.abs().max())
            res[name] = {"ms": round(ms, 4), "pairs_per_s": pairs / ms * 1e3, "max_diff_vs_window": dmax}
        except Exception as e:  # noqa: BLE001
            res[name] = {"error": str(e)[:200]}
    setenv()
    print(json.dumps({"S": S, "M": M, "potential": potential, "results": res}), flush=True)


if __name__ == "__main__":
    shapes = [(500, 524288), (500, 1 << 20), (5000, 524288), (1000, 250000), (300, 300000), (2000, 65536),
              (20000, 200000), (65536, 65536), (262144, 262144), (150, 11031), (4096, 4096), (1 << 20, 1000)]
    if len(sys.argv) > 1 and sys.argv[1] == "big":
        shapes = [(1 << 20, 1 << 20)]
    if len(sys.argv) > 1 and sys.argv[1] == "multi":
        shapes = [(2000, 65536), (5000, 524288), (20000, 200000), (65536, 65536), (262144, 262144)]
    pot_shapes = []
    if len(sys.argv) > 1 and "x" in sys.argv[1]:                 # 500x524288,p65536x4194304 (p: potential)
        toks = sys.argv[1].split(",")
        shapes = [tuple(int(v) for v in t.split("x")) for t in toks if not t.startswith("p")]
        pot_shapes = [tuple(int(v) for v in t[1:].split("x")) for t in toks if t.startswith("p")]
    for S, M in shapes:
        case(S, M)
    for S, M in pot_shapes:
        case(S, M, potential=True)
    if len(sys.argv) <= 1 or sys.argv[1] in ("big",):
        for S, M in ([(1 << 20, 1 << 20)] if len(sys.argv) > 1 else [(65536, 65536), (1 << 20, 1000)]):
            case(S, M, potential=True)
