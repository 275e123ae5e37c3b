"""Persistent cooperative loop vs device-driven launches across cloud sizes (synthetic noisy sphere, spatially coherent
patches of ~`rows` points): where does the persistent kernel stop winning?  CUDA events around the native call."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200"))
    import torch, dnp_native as dn
    n, rows, diffuse = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(3)
    d = torch.randn(n, 3, generator=g)
    d = d / d.norm(dim=-1, keepdim=True)
    pc = torch.cat([d * 0.45 + 0.002 * torch.randn(n, 3, generator=g), d], 1).float()
    key = ((pc[:, :3] + 0.6) * 16).long().clamp(0, 31)
    order = torch.argsort(key[:, 0] * 1024 + key[:, 1] * 32 + key[:, 2], stable=True)
    P = max(n // rows, 1)
    chunks = torch.chunk(order, P)
    offsets = torch.tensor([0] + [c.numel() for c in chunks]).cumsum(0)
    perm = torch.cat(chunks).to(dev)
    pc = pc.to(dev)
    best = 1e9
    for _ in range(3):
        work = pc.clone()
        E = torch.zeros(n, 3, device=dev)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        dn.patch_propagate(work, perm, offsets, 0, E, diffuse=bool(diffuse), eps=1e-5)
        b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    print(json.dumps({"ms": best, "P": len(chunks), "us_per_step": best * 1e3 / len(chunks)}))
    sys.exit(0)
for n, rows in ((11000, 150), (30000, 300), (100000, 500), (100000, 1000), (200000, 500), (300000, 600), (300000, 1000)):
    for diffuse in (0, 1):
        line = f"N={n} rows/patch={rows} diffuse={diffuse}:"
        for mode in ("persistent", "device"):
            env = dict(os.environ, DNP_PATCH_LOOP=mode, DNP_LOOP_MAX_PAIRS="1e12")
            r = subprocess.run([sys.executable, __file__, "child", str(n), str(rows), str(diffuse)], env=env, capture_output=True, text=True)
            try:
                dd = json.loads(r.stdout.strip().splitlines()[-1])
                line += f"  {mode} {dd['ms']:.2f} ms ({dd['us_per_step']:.1f} us/step, P={dd['P']})"
            except Exception:
                line += f"  {mode} failed: {r.stderr.strip().splitlines()[-1][:120] if r.stderr.strip() else '?'}"
        print(line, flush=True)
