"""Device timings of the text IO kernels on a 1M-point cloud (export -> parse), for profiles/."""
import json, os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200")); sys.path.insert(0, ROOT)
import dnp_native as dn, dnp_util
import bench
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(8)
rows = torch.randn(1 << 20, 6, generator=g).to(dev)
text = dnp_util.export_text(rows.transpose(0, 1))
txt_dev = torch.frombuffer(bytearray(text), dtype=torch.uint8).to(dev)
def best(fn, reps=3):
    fn(); torch.cuda.synchronize(); b = 1e9
    for _ in range(reps):
        a, c = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); c.record(); torch.cuda.synchronize(); b = min(b, a.elapsed_time(c))
    return b
n_lines = text.count(b"\n") + 1
out = {"text_bytes": len(text),
       "format_ms": best(lambda: bench._format_only(dn, rows)),
       "parse_ms": best(lambda: bench._parse_only(dn, txt_dev, n_lines))}
short = "\n".join(" ".join("%.6f" % v for v in r) for r in rows[:200000].cpu().tolist()).encode()
sd = torch.frombuffer(bytearray(short), dtype=torch.uint8).to(dev)
out["parse_200k_pts_6_decimals_ms"] = best(lambda: bench._parse_only(dn, sd, 200000))
out["parse_200k_text_bytes"] = len(short)
out["parse_gbs"] = len(text) / out["parse_ms"] / 1e6
out["format_gbs_written"] = len(text) / out["format_ms"] / 1e6
back, mask, bad, lines = dn.xyz_parse(text)
out["round_trip_bit_exact"] = bool(torch.equal(back.view(torch.int32), rows.view(torch.int32)))
print(json.dumps(out))
