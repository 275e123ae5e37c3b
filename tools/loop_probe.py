"""The fandisk patch loop (tests/golden/patch_prop.npz, 72 patches of 100-268 rows, 11 031 points) under the
schedules / plans of the greedy driver: wall time per run (CUDA events) for each override set."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200"))
    import numpy as np, torch, field_utils
    g = np.load(os.path.join(ROOT, "tests", "golden", "patch_prop.npz"))
    dev = torch.device("cuda:0")
    pc = torch.from_numpy(g["pc_in"]).to(dev)
    patches = [torch.from_numpy(g["perm"][g["offsets"][i]:g["offsets"][i + 1]]).to(dev) for i in range(int(g["P"]))]
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    best = 1e9
    for _ in range(reps):
        work = pc.clone()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        field_utils.strongest_field_propagation(work, [], patches)
        b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    # the native call alone (CSR prebuilt, start patch given): what the schedule itself costs
    import dnp_native as dn
    perm = torch.from_numpy(g["perm"]).to(dev)
    offsets = torch.from_numpy(g["offsets"]).to(dev)
    best_call = 1e9
    for _ in range(reps):
        work = pc.clone()
        E = torch.zeros(pc.shape[0], 3, device=dev)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        dn.patch_propagate(work, perm, offsets, 0, E, diffuse=False, eps=1e-5)
        b.record(); torch.cuda.synchronize()
        best_call = min(best_call, a.elapsed_time(b))
    print(json.dumps({"ms": best, "native_call_ms": best_call}))
    sys.exit(0)
for name, over in (("auto", {}), ("tile256", {"DNP_FORCE_TILE": "256"}), ("tile64 nsplit1", {"DNP_FORCE_TILE": "64", "DNP_FORCE_NSPLIT": "1"}),
                   ("tile64 nsplit3", {"DNP_FORCE_TILE": "64", "DNP_FORCE_NSPLIT": "3"}),
                   ("host loop", {"DNP_PATCH_LOOP": "host"})):
    r = subprocess.run([sys.executable, __file__, "child"], env=dict(os.environ, **over), capture_output=True, text=True)
    print(name, r.stdout.strip().splitlines()[-1] if r.stdout.strip() else r.stderr[-300:], flush=True)
