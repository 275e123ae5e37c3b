"""CPU timing of the reference-style text IO (oracle restatement = the reference's per-line Python loops)."""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as orc
rng = np.random.default_rng(0)
pc = rng.normal(size=(50000, 6)).astype(np.float32)
t0 = time.perf_counter(); txt = orc.export_text(pc); t1 = time.perf_counter()
back = orc.xyz2tensor(txt); t2 = time.perf_counter()
print(f"export {1e6*(t1-t0)/pc.size:.2f} us/value, parse {1e6*(t2-t1)/pc.size:.2f} us/value, bytes/value {len(txt)/pc.size:.1f}")
