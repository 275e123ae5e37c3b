"""Small instance of every launch shape of the hot path in one process, each checked against the oracle where that is
cheap: the window kernels are forced at small sizes through DNP_RING=2 (one launch with the fused fix-up, shares on
two lanes, whole tiles on 2 and 4 lanes), the four patch-loop schedules, the point propagation, partition and text
IO.  (Written as the workload of a compute-sanitizer pass; that tool is closed on this GPU pool, so this is run plain:
`python tools/small_shapes.py all`.)  Development tool."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200"))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import dnp_native as dn  # noqa: E402
import dnp_util  # noqa: E402
import field_utils  # noqa: E402
from oracle import oracle as orc  # noqa: E402

dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(3)


def cloud(n):
    d = torch.nn.functional.normalize(torch.randn(n, 3, generator=g), dim=1)
    pc = torch.cat([d * 0.45 + torch.randn(n, 3, generator=g) * 0.01, d], 1).float()
    pc[torch.rand(n, generator=g) < 0.5, 3:] *= -1
    return pc


def close(a, b, what):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert np.abs(a - b).max() <= 2e-4 * np.abs(b).max(), what
    print("ok", what, flush=True)


def leaf(tag, S, M, **env):
    for k in list(os.environ):
        if k.startswith("DNP_RING"):
            del os.environ[k]
    os.environ.update({k: str(v) for k, v in env.items()})
    src, tgt = cloud(S), cloud(M)
    src[1, :3] = tgt[5, :3]                                    # a coincident pair: the safe re-run of a fold
    E = field_utils.field_grad(src.to(dev), tgt.to(dev)).cpu().numpy()
    phi = field_utils.potential(src.to(dev), tgt.to(dev)).cpu().numpy()
    close(E, orc.field_grad(src.numpy(), tgt.numpy()), f"field {tag} {S}x{M}")
    keep = np.arange(M) != 5        # the reference zeroes that target's whole leaf total, the kernels drop the pair
    close(phi[keep], orc.potential(src.numpy(), tgt.numpy())[keep], f"potential {tag} {S}x{M}")


which = sys.argv[1] if len(sys.argv) > 1 else "all"
if which in ("all", "leaf"):
    leaf("shared-memory kernel", 700, 3001)
    leaf("window, one launch (fused fix-up), 8 targets/thread", 600, 5000, DNP_RING=2, DNP_RING_TP=4)
    leaf("window, one launch, 2 targets/thread", 600, 3000, DNP_RING=2, DNP_RING_TP=1)
    leaf("window, shares on 2 lanes + reduction", 2100, 5000, DNP_RING=2, DNP_RING_MODE=1)
    leaf("window, whole tiles, 2 lanes", 2100, 5000, DNP_RING=2, DNP_RING_MODE=2, DNP_RING_LANES=2)
    leaf("window, whole tiles, 4 lanes", 2100, 5000, DNP_RING=2, DNP_RING_MODE=2, DNP_RING_LANES=4)
    for k in list(os.environ):
        if k.startswith("DNP_RING"):
            del os.environ[k]
if which in ("all", "scores"):
    pc = cloud(6000)
    m = field_utils.measure_mean_potential(pc.to(dev)).item()
    close(m, orc.measure_mean_potential(pc.numpy()), "lattice mean")
    a, b = cloud(900), cloud(1100)
    w, _ = field_utils.field_edge_calculator(a.to(dev), b.to(dev))
    close(float(w), orc.field_edge_calculator(a.numpy(), b.numpy())[0], "edge score")
    ref, tgt = cloud(1500), cloud(2500)
    out = field_utils.reference_field(ref.to(dev), tgt.clone().to(dev))
    assert out.shape == (2500, 6)
    print("ok reference_field", flush=True)
if which in ("all", "drivers"):
    pts = cloud(1200)
    want, _, _ = orc.points_propagate(pts.numpy(), diffuse=True)
    work = pts.clone().to(dev)
    field_utils.strongest_field_propagation_points(work, diffuse=True)
    assert np.array_equal(work.cpu().numpy(), want)
    print("ok point propagation (cluster kernel)", flush=True)
    pc = cloud(4000)
    order = torch.argsort(pc[:, 0] * 5 + pc[:, 1], stable=True)
    patches = [c.to(dev) for c in torch.chunk(order, 12)]
    results = []
    for mode in ("persistent", "slab", "device", "host"):
        os.environ["DNP_PATCH_LOOP"] = mode
        work = pc.clone().to(dev)
        field_utils.strongest_field_propagation(work, [], patches, diffuse=True)
        results.append(work.cpu())
    del os.environ["DNP_PATCH_LOOP"]
    assert all(torch.equal(torch.sign(results[0][:, 3:] * pc[:, 3:]).sum(1), torch.sign(r[:, 3:] * pc[:, 3:]).sum(1))
               for r in results[1:])
    print("ok patch propagation, three schedules agree", flush=True)
    reps = [(p[:40], p[40:]) for p in patches]
    work = pc.clone().to(dev)
    field_utils.strongest_field_propagation_reps(work, reps, diffuse=True)
    print("ok reps propagation", flush=True)
if which in ("all", "next"):
    pc = cloud(20000).to(dev)
    parts = dnp_util.divide_pc(pc, 6, min_patch=50)
    print("ok partition:", len(parts), "patches", flush=True)
    txt = dnp_util.export_text(pc[:3000].transpose(0, 1))
    back = dnp_util.xyz2tensor_cuda(txt)
    assert torch.equal(back.view(torch.int32), pc[:3000].view(torch.int32))
    print("ok text round trip", flush=True)
torch.cuda.synchronize()
print("workload done:", dn.launch_count(), "libdipole launches", flush=True)
