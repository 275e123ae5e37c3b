"""Persistent cooperative loop vs slab schedule vs device-driven launches across cloud sizes (synthetic noisy sphere,
spatially coherent patches of ~`rows` points), one process: which schedule should a mid-size loop take?  CUDA events
around the native call, best of 3; visit orders compared with the device-driven launches."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200"))
import torch, dnp_native as dn

dev = torch.device("cuda:0")
SIZES = [(60000, 500), (100000, 500), (100000, 1000), (150000, 500), (200000, 500), (200000, 1000), (300000, 600),
         (300000, 1000), (400000, 500)]
for n, rows in SIZES:
    g = torch.Generator().manual_seed(3)
    d = torch.randn(n, 3, generator=g)
    d = d / d.norm(dim=-1, keepdim=True)
    pc = torch.cat([d * 0.45 + 0.002 * torch.randn(n, 3, generator=g), d], 1).float()
    key = ((pc[:, :3] + 0.6) * 16).long().clamp(0, 31)
    order = torch.argsort(key[:, 0] * 1024 + key[:, 1] * 32 + key[:, 2], stable=True)
    chunks = torch.chunk(order, max(n // rows, 1))
    offsets = torch.tensor([0] + [c.numel() for c in chunks]).cumsum(0)
    perm = torch.cat(chunks).to(dev)
    pcd = pc.to(dev)
    for diffuse in (False, True):
        rec = {"N": n, "rows_per_patch": rows, "P": len(chunks), "diffuse": diffuse}
        orders = {}
        for mode in ("persistent", "slab", "device"):
            os.environ["DNP_PATCH_LOOP"] = mode
            os.environ["DNP_LOOP_MAX_PAIRS"] = "1e12"
            best = 1e9
            try:
                for it in range(3):
                    work, E = pcd.clone(), torch.zeros(n, 3, device=dev)
                    torch.cuda.synchronize()
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record()
                    tr = dn.patch_propagate(work, perm, offsets, 0, E, diffuse=diffuse, eps=1e-5, want_trace=(it == 0))
                    b.record(); torch.cuda.synchronize()
                    if it == 0:
                        orders[mode] = list(tr[0])
                    else:
                        best = min(best, a.elapsed_time(b))
                rec[mode + "_ms"] = round(best, 3)
            except dn.DipoleError as e:
                rec[mode + "_ms"] = None
        rec["same_order"] = all(orders.get(m) == orders.get("device") for m in orders)
        print(json.dumps(rec), flush=True)
