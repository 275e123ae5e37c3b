"""Host-side profile of strongest_field_propagation_reps on the cfg3 stand-in (1M points, 1000 patches, 500
representatives each): where the ~35 ms outside the two field stages go."""
import cProfile, os, pstats, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200"))
import torch, field_utils
from bench import surface_cloud
dev = torch.device("cuda:0")
pc3 = surface_cloud(1 << 20, 10)
key = ((pc3[:, :3] + 0.6) * 8).long().clamp(0, 15)
order = torch.argsort(key[:, 0] * 256 + key[:, 1] * 16 + key[:, 2], stable=True)
chunks = torch.chunk(order, 1000)
reps = [(c[:500].to(dev), c[500:].to(dev)) for c in chunks]
pc3 = pc3.to(dev)
for _ in range(2):
    w = pc3.clone(); field_utils.strongest_field_propagation_reps(w, reps, diffuse=True); torch.cuda.synchronize()
w = pc3.clone(); torch.cuda.synchronize()
t0 = time.perf_counter()
pr = cProfile.Profile(); pr.enable()
field_utils.strongest_field_propagation_reps(w, reps, diffuse=True)
torch.cuda.synchronize()
pr.disable()
print(f"total {1e3 * (time.perf_counter() - t0):.1f} ms")
pstats.Stats(pr).sort_stats("cumulative").print_stats(22)
