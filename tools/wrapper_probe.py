"""Host cost of one small call through the Python mirror (ctypes -> C ABI): cProfile of 5000 field_grad calls on a
100-source x 2000-target problem (GPU time ~10 us, so the loop is host bound)."""
import cProfile, os, pstats, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200"))
import torch, dnp_native as dn, field_utils as fu
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
src = torch.randn(100, 6, generator=g).to(dev)
tgt = torch.randn(2000, 6, generator=g).to(dev)
for _ in range(100): fu.field_grad(src, tgt)
torch.cuda.synchronize()
n = 5000
t0 = time.perf_counter()
for _ in range(n): fu.field_grad(src, tgt)
torch.cuda.synchronize()
print(f"field_grad small call: {(time.perf_counter() - t0) / n * 1e6:.1f} us per call")
t0 = time.perf_counter()
for _ in range(n): fu.potential(src, tgt)
torch.cuda.synchronize()
print(f"potential small call: {(time.perf_counter() - t0) / n * 1e6:.1f} us per call")
pr = cProfile.Profile(); pr.enable()
for _ in range(n): fu.field_grad(src, tgt)
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(14)
