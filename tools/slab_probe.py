"""Config-3 stand-in (1M points, 1000 patches, 500 representatives each, diffuse) through the greedy loop's schedules:
the slab schedule (dnp_slab.cu, default at this size) against the device-driven launches.  CUDA events around the
native loop call and around the whole strongest_field_propagation_reps call; decisions compared."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200"))
import torch
import dnp_native as dn
import field_utils
from bench import surface_cloud


def ev_ms(fn, reps=3):
    best = float("inf")
    out = None
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        out = fn()
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best, out


def main():
    dev = torch.device("cuda:0")
    n_pts = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
    P = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
    n_rep = int(sys.argv[3]) if len(sys.argv) > 3 else 500
    pc = surface_cloud(n_pts, 10)
    key = ((pc[:, :3] + 0.6) * 8).long().clamp(0, 15)
    order = torch.argsort(key[:, 0] * 256 + key[:, 1] * 16 + key[:, 2], stable=True)
    chunks = torch.chunk(order, P)
    reps = [(c[:n_rep].to(dev), c[n_rep:].to(dev)) for c in chunks]
    perm = torch.cat([r for r, _ in reps])
    offsets = torch.tensor([0] + [r.numel() for r, _ in reps], dtype=torch.int64).cumsum(0)
    rest = torch.cat([q for _, q in reps])
    rest_off = torch.tensor([0] + [q.numel() for _, q in reps], dtype=torch.int64).cumsum(0)
    pcd = pc.to(dev)
    res = {}
    for diffuse in (True, False):
        for mode in (("slab",) if os.environ.get("SLAB_ONLY") else ("slab", "slab2", "slab4", "device")):
            os.environ["DNP_PATCH_LOOP"] = mode[:4] if mode.startswith("slab") else mode
            os.environ.pop("DNP_RING_LANES", None)
            if mode in ("slab2", "slab4"):
                os.environ["DNP_RING_LANES"] = mode[-1]

            def loop():
                work, E = pcd.clone(), torch.zeros(n_pts, 3, device=dev)
                return dn.patch_propagate(work, perm, offsets, 0, E, diffuse=diffuse, rest_perm=rest,
                                          rest_offsets=rest_off, tgt_idx=perm, want_trace=True), work, E

            def whole():
                work = pcd.clone()
                field_utils.strongest_field_propagation_reps(work, reps, diffuse=diffuse)
                return work
            loop()
            n0 = dn.launch_count()
            ms_loop, ((o, v), work, E) = ev_ms(loop)
            launches = (dn.launch_count() - n0) // 3
            ms_whole, w2 = ev_ms(whole, 2)
            dn.profile_begin()
            loop()
            torch.cuda.synchronize()
            prof = dn.profile_end()
            res[(diffuse, mode)] = (list(o), work, E, w2)
            pairs = float(sum(r.numel() for r, _ in reps)) * perm.numel()
            print(json.dumps({"points": n_pts, "patches": P, "reps": n_rep, "diffuse": diffuse, "schedule": mode,
                              "loop_ms": round(ms_loop, 2), "loop_pairs_per_s": pairs / (ms_loop * 1e-3),
                              "whole_call_ms": round(ms_whole, 2), "launches": int(launches),
                              "lanes": os.environ.get("DNP_RING_LANES", "auto"),
                              "field_launches_ms": [round(float(x), 2) for x in prof][:4]}), flush=True)
        if os.environ.get("SLAB_ONLY"):
            break
        a, b = res[(diffuse, "slab")], res[(diffuse, "device")]
        same_order = a[0] == b[0]
        first = next((i for i, (x, y) in enumerate(zip(a[0], b[0])) if x != y), -1)
        scale = b[2].norm(dim=-1).max()
        print(json.dumps({"diffuse": diffuse, "same_visit_order": same_order, "first_difference": first,
                          "same_normals_after_loop": bool(torch.equal(a[1], b[1])),
                          "max_rel_dE": ((a[2] - b[2]).norm(dim=-1).max() / scale).item(),
                          "final_normals_differ": int((a[3] != b[3]).any(dim=1).sum().item())}), flush=True)
    os.environ.pop("DNP_PATCH_LOOP", None)


if __name__ == "__main__":
    main()
