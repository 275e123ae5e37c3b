#!/usr/bin/env python
"""Benchmark of the dipole-field hot path (BASELINE.json metric: dipole pair-interactions/sec).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--points P]

A "step" is one field_grad pass (field_utils.py:61-116) of S = P replicated sources over this GPU's shard
of M = P targets (BASELINE config 5, the synthetic all-pairs sweep at 1M x 1M per GPU).  Multi-GPU runs
shard the TARGETS over the ranks with the sources replicated and no data-path collective (weak scaling:
the global target set grows with N).  One JSON line is printed by rank 0.

  value        pairs/s, inputs resident in HBM, K steps between barrier+synchronize, max over ranks
  e2e          the same metric through the reference-facing call field_utils.field_grad(host tensors):
               pinned host buffers in, host tensor out, H2D/D2H copies inside the timed region
  roofline     the all-pairs kernel against the FP32 FMA pipe (35 flop/pair); the kernel
               durations come from CUDA events recorded by the library around each launch in the timed region
  cpu_baseline the torch port of the reference's CPU path (oracle/torch_port.py) on the box's host cores,
               bounded sample (rank 0, N = 1 only)

`--impl reference` times that CPU port on the same metric and config (bounded sample per step).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "dipole-normal-prop_b200"))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "dipole_pair_interactions_per_sec"
UNIT = "pairs/s"
FLOP_PER_PAIR_FIELD = 35.0      # 23 FP32-pipe instructions + 2 SFU ops
FLOP_PER_PAIR_POTENTIAL = 18.0
EPS = 1e-5


def cloud(n, seed, cols=6, pin=False):
    """Seeded synthetic cloud: positions uniform in [-0.5, 0.5]^3, unit moments."""
    g = torch.Generator().manual_seed(seed)
    xyz = torch.rand(n, 3, generator=g) - 0.5
    if cols == 6:
        nr = torch.randn(n, 3, generator=g)
        xyz = torch.cat([xyz, nr / nr.norm(dim=-1, keepdim=True)], 1)
    out = xyz.float().contiguous()
    return out.pin_memory() if pin else out


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                 "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *exc):
        if self.proc:
            time.sleep(0.25)
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        mhz, mx, reasons, power = [], [], set(), []
        for r in self.rows:
            try:
                mhz.append(float(r[0])); mx.append(float(r[1])); power.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not mhz:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        mhz.sort()
        return {"sm_mhz": mhz[len(mhz) // 2], "sm_min_mhz": mhz[0], "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(mhz), "power_w_max": max(power)}


# ---------------------------------------------------------------------------------------------- CPU arm
def time_cpu_port(n_src, n_tgt, repeats=1, potential=False):
    """Seconds per field_grad (or potential) of n_src x n_tgt points in the torch port on all host cores (best of
    `repeats`)."""
    from oracle import torch_port
    src, tgt = cloud(n_src, 1), cloud(n_tgt, 2, cols=3)
    best = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        if potential:
            torch_port.potential(src, tgt)
        else:
            torch_port.field_grad(src, tgt, eps=EPS)
        best = min(best, time.perf_counter() - t0)
    return best


def cpu_baseline_block(n):
    torch.set_num_threads(os.cpu_count() or 1)
    time_cpu_port(2048, 2048)                          # warm the thread pool and the allocator
    m = 2 * n                                          # > max_pts = 15000 at the default n: the chunking runs too
    sec = time_cpu_port(n, m)
    sec_phi = time_cpu_port(n, m, potential=True)
    return {"value": n * m / sec, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
            "potential_pairs_per_s": n * m / sec_phi,
            "host_cpus": os.cpu_count(), "drivers": cpu_driver_seconds(full=False),
            "sample": f"oracle/torch_port.field_grad (the reference's op chain and chunking, field_utils.py:61-116), "
                      f"float32, {n} sources x {m} targets = {n * m:.3e} pairs, one run of {sec:.1f} s"}


def _golden(name):
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))


def _csr_patches(perm, off, dev=None):
    out = [torch.from_numpy(perm[off[i]:off[i + 1]].copy()) for i in range(len(off) - 1)]
    return [p.to(dev) for p in out] if dev is not None else out


def cpu_driver_seconds(full):
    """Wall seconds of the reference's DRIVERS on this box's host cores (BASELINE.json metric "e2e orient sec";
    BASELINE.md section 4 item 4): oracle/torch_drivers.py = the reference's op sequence (bit-equal outputs, pinned
    in tests/test_oracle_drivers.py), stage names as in the reference's MyTimer blocks ('propagating field' /
    'fix global orientation' / 'calculating field').  full=False: the two cheap configurations only."""
    from oracle import torch_drivers as td
    torch.set_num_threads(os.cpu_count() or 1)
    out, stages = {}, {}

    def timed(key, fn):
        t0 = time.perf_counter()
        fn()
        out[key] = time.perf_counter() - t0

    def pipeline(key, propagate, pts):
        t0 = time.perf_counter()
        propagate()
        t1 = time.perf_counter()
        flip = bool(td.lattice_mean_potential(pts) < 0)
        if flip:
            pts[:, 3:] *= -1
        t2 = time.perf_counter()
        out[key] = t2 - t0
        stages[key] = {"propagating field": t1 - t0, "fix global orientation": t2 - t1}

    g = _golden("points_prop")
    pts = torch.from_numpy(g["pc_in"][:2000].copy())
    pipeline("cfg1_ok_xyz_first_2000_pts_e2e_s", lambda: td.propagate_points(pts, diffuse=True), pts)
    g2 = _golden("patch_prop")
    groups = _csr_patches(g2["perm"], g2["offsets"])
    pts2 = torch.from_numpy(g2["pc_in"].copy())
    pipeline("cfg2_fandisk_72_patches_e2e_s", lambda: td.propagate_patches(pts2, [], groups), pts2)
    if full:
        pts1 = torch.from_numpy(g["pc_in"].copy())
        pipeline("cfg1_ok_xyz_10000_pts_e2e_s", lambda: td.propagate_points(pts1), pts1)
        pc3, reps = cfg3_standin()
        st = {}
        pipeline("cfg3_reps_60000_pts_48_patches_e2e_s",
                 lambda: td.propagate_representatives(pc3, reps, diffuse=True, stage_seconds=st), pc3)
        stages["cfg3_reps_60000_pts_48_patches_e2e_s"].update({"propagation (greedy loop)": st.get("propagation"),
                                                               "final field reps -> rest": st.get("final_field")})
        ref, tgt = cfg4_standin()
        timed("cfg4_transfer_20000_x_40000_e2e_s", lambda: td.transfer_orientation(ref, tgt))
        stages["cfg4_transfer_20000_x_40000_e2e_s"] = {"calculating field": out["cfg4_transfer_20000_x_40000_e2e_s"]}
    out["stages"] = stages
    out["cores"] = torch.get_num_threads()
    return out


# ---------------------------------------------------------------------------------------------- multi-GPU product path
def dist_parity(dn, dist, dnp_dist, dev, rank, world):
    """The assertions of tests/test_gpu_dist.py on this run's ranks, before anything is timed: every sharded call
    of dnp_dist against the same call on one GPU (each rank computes the unsharded answer itself)."""
    src, tgt = cloud(20001, 1).to(dev), cloud(150003, 2).to(dev)          # uneven shards
    checks = {}
    E = dn.field_grad(src, tgt)
    mag = E.norm(dim=-1)
    E_full = dnp_dist.field_grad_sharded(src, tgt, gather=True)
    checks["field_grad_gathered_within_tolerance"] = bool(((E_full - E).norm(dim=-1) <= 1e-4 * mag + 1e-6 * mag.max()).all())
    a, b = dnp_dist.shard_range(tgt.shape[0], rank, world)
    checks["local_block_is_the_gathered_slice"] = bool(torch.equal(dnp_dist.field_grad_sharded(src, tgt), E_full[a:b]))
    phi = dn.potential(src, tgt)
    phi_full = dnp_dist.potential_sharded(src, tgt, gather=True)
    checks["potential_gathered_within_tolerance"] = bool(((phi_full - phi).abs() <= 1e-4 * phi.abs() + 1e-5 * phi.abs().max()).all())
    mean = dnp_dist.measure_mean_potential_sharded(src)
    want = dn.lattice_mean_potential(src)
    checks["lattice_mean_sources_sharded"] = bool((mean - want).abs() <= 1e-4 * want.abs())
    perm = torch.randperm(20001, generator=torch.Generator().manual_seed(3)).to(dev)
    offsets = torch.tensor([0, 400, 401, 1200, 2500, 3010, 20001])
    edges = torch.tensor([[0, 1], [1, 0], [2, 3], [3, 2], [4, 0], [1, 1], [5, 4]])
    inter = dnp_dist.edge_interactions_sharded(src, perm, offsets, edges)
    one = dn.edge_interactions_batched(src, perm, offsets.to(dev), edges.to(dev))
    checks["edge_interactions_allreduced_within_tolerance"] = bool(((inter - one).abs() <= 2e-4 * one.abs() + 1e-3).all())
    # the C ABI's own multi-GPU entry points (dnp_comm.cu: NCCL enqueued by the library on the kernels' stream; the
    # 128-byte id is the only thing that travels through torch.distributed)
    ident = (dn.comm_unique_id() if rank == 0 else torch.zeros(128, dtype=torch.uint8)).to(dev)
    dist.broadcast(ident, 0)
    dn.comm_init(ident.cpu(), rank, world)
    try:
        checks["c_abi_field_grad_sharded_equals_dnp_dist"] = bool(torch.equal(dn.field_grad_sharded(src, tgt, gather=True), E_full))
        abi_inter = dn.edge_interactions_sharded(src, perm, offsets, edges)
        checks["c_abi_edge_interactions_allreduced"] = bool(((abi_inter - one).abs() <= 2e-4 * one.abs() + 1e-3).all())
        abi_mean = dn.lattice_mean_potential_sharded(src)
        checks["c_abi_lattice_mean_sharded"] = bool((abi_mean - want).abs() <= 1e-4 * want.abs())
        torch.cuda.synchronize()
    finally:
        dn.comm_destroy()
    # identical on every rank (all_gather / all_reduce results): compare bit patterns through a MAX and a MIN reduce
    sig = torch.stack([E_full.view(torch.int32).sum(dtype=torch.int64), inter.view(torch.int32).sum(dtype=torch.int64)])
    hi, lo = sig.clone(), sig.clone()
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    checks["results_identical_on_all_ranks"] = bool(torch.equal(hi, lo))
    ok = torch.tensor([int(all(checks.values()))], device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    return {"ok": bool(ok.item()), "world": world, "rank0_checks": checks}


def multi_gpu_blocks(args, dn, dist, dev, rank, world):
    """Strong scaling of the product's multi-GPU layer (dnp_dist.py): a FIXED global problem split over the ranks,
    starting from rank 0's pinned host buffers and ending with the full result in rank 0's host memory.
      field: G sources x G targets; broadcast of the sources (24 B each) and the targets (12 B each), each rank's
             target shard through field_grad, all_gather of E (12 B per target), device -> host on rank 0
      edges: flip scores of a 1000-patch graph of the 1M-point cloud (field_utils.py:145-160); each rank sums over
             its slice of every target patch, ONE all_reduce(SUM) over the edge vector combines them."""
    import dnp_dist

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def ev_ms(fn):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sync_all()
        a.record()
        r = fn()
        b.record()
        sync_all()
        return max_over_ranks(a.elapsed_time(b)), r

    G = args.strong_points
    out = {}
    src_host = cloud(G, 11, pin=True) if rank == 0 else None
    tgt_host = cloud(G, 12, cols=3, pin=True) if rank == 0 else None
    E_host = torch.empty(G, 3).pin_memory() if rank == 0 else None
    src = torch.empty(G, 6, device=dev)
    tgt = torch.empty(G, 3, device=dev)

    def field_step():
        if rank == 0:
            src.copy_(src_host, non_blocking=True)
            tgt.copy_(tgt_host, non_blocking=True)
        if world > 1:
            dist.broadcast(src, 0)
            dist.broadcast(tgt, 0)
        E = dnp_dist.field_grad_sharded(src, tgt, eps=EPS, gather=True)
        if rank == 0:
            E_host.copy_(E, non_blocking=True)
        return E

    field_step()
    times = [ev_ms(field_step)[0] for _ in range(2)]
    ms = min(times)
    compute_ms, _ = ev_ms(lambda: dnp_dist.field_grad_sharded(src, tgt, eps=EPS))
    bcast_ms = ev_ms(lambda: (dist.broadcast(src, 0), dist.broadcast(tgt, 0)))[0] if world > 1 else 0.0
    a, b = dnp_dist.shard_range(G, rank, world)
    local = torch.empty(b - a, 3, device=dev)
    gather_ms = ev_ms(lambda: dnp_dist._gather_rows(local, G, None))[0] if world > 1 else 0.0
    out["field"] = {
        "sources": G, "targets": G, "ms_per_step": ms, "pairs_per_s": float(G) * G / (ms * 1e-3),
        "shard_compute_ms": compute_ms, "broadcast_ms": bcast_ms, "all_gather_ms": gather_ms,
        "broadcast_bytes": G * 36, "all_gather_bytes": G * 12, "h2d_bytes": G * 36, "d2h_bytes": G * 12,
        "bound_by": "shard compute" if compute_ms >= max(bcast_ms, gather_ms) else "collective",
        "api": "dnp_dist.field_grad_sharded(gather=True): targets block-sharded, sources replicated by broadcast"}
    del src, tgt, local

    # ---- flip scores of a patch graph, one all_reduce
    pc = surface_cloud(1 << 20, 10)
    key = ((pc[:, :3] + 0.6) * 8).long().clamp(0, 15)
    order = torch.argsort(key[:, 0] * 256 + key[:, 1] * 16 + key[:, 2], stable=True)
    chunks = torch.chunk(order, 1000)
    perm = torch.cat(chunks).to(dev)
    offsets = torch.zeros(1001, dtype=torch.int64)
    offsets[1:] = torch.tensor([c.numel() for c in chunks]).cumsum(0)
    e = [(i, (i + d) % 1000) for i in range(1000) for d in (1, 2, 3)]
    edges = torch.tensor(e + [(j, i) for i, j in e])
    pcd = pc.to(dev)
    dnp_dist.edge_scores_sharded(pcd, perm, offsets, edges)
    ms_e, (w, _) = ev_ms(lambda: dnp_dist.edge_scores_sharded(pcd, perm, offsets, edges))
    vec = torch.zeros(edges.shape[0], device=dev)
    ar_ms = ev_ms(lambda: dist.all_reduce(vec))[0] if world > 1 else 0.0
    pairs = float(sum(chunks[i].numel() * chunks[j].numel() for i, j in edges.tolist()))
    out["edge_scores"] = {"patches": 1000, "edges": int(edges.shape[0]), "pairs": pairs, "ms": ms_e,
                          "pairs_per_s": pairs / (ms_e * 1e-3), "all_reduce_ms": ar_ms,
                          "all_reduce_bytes": int(edges.shape[0]) * 4, "checksum": float(w.double().abs().sum().item()),
                          "api": "dnp_dist.edge_scores_sharded: per-patch target slices + one all_reduce(SUM)"}

    # ---- BASELINE config 3 stand-in on the ranks: greedy propagation over the representatives of 1000 patches of the
    # same 1M-point cloud, targets sharded by patch ownership with one all-reduce of the interaction vector per step
    # (dnp_patch_propagate_sharded_f32), then the reps -> rest field sharded over the remaining rows
    reps = [(c[:500].to(dev), c[500:].to(dev)) for c in chunks]
    work = pcd.clone()
    dnp_dist.strongest_field_propagation_reps_sharded(work, reps, diffuse=True)

    def run_reps():
        w2 = pcd.clone()
        dnp_dist.strongest_field_propagation_reps_sharded(w2, reps, diffuse=True)
        return w2
    ms_r, w2 = ev_ms(run_reps)
    sig = w2[:, 3:].contiguous().view(torch.int32).sum(dtype=torch.int64).reshape(1)
    hi, lo = sig.clone(), sig.clone()
    if world > 1:
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    flipped = int((w2[:, 3:] != pcd[:, 3:]).any(dim=1).sum().item())
    out["cfg3_reps_propagation"] = {
        "points": 1 << 20, "patches": 1000, "representatives_per_patch": 500, "diffuse": True, "ms": ms_r,
        "identical_on_all_ranks": bool(torch.equal(hi, lo)), "points_flipped": flipped,
        "schedule": "slab (dnp_slab.cu): every rank computes the field slabs of the patches it owns against all "
                    "representatives, then the greedy decisions run on the all-reduced P x P matrix",
        "all_reduce_bytes": [1000 * 1000 * 4, 500 * 1000 * 12], "steps": 999,
        "api": "dnp_dist.strongest_field_propagation_reps_sharded: patches owned round robin, one all_reduce(SUM) "
               "of the [P, P] interaction matrix and one of the [T, 3] field, final field sharded over the remaining rows"}
    dnp_dist.release_comm()
    return out


def gpu_driver_seconds(field_utils, dev):
    """The same five driver workloads as cpu_driver_seconds through the drop-in API, HOST tensors in and out
    (device copies inside the timed region), wall clock, best of 2; each result is compared with the reference's
    recorded decisions (tests/golden, generated by the unmodified reference)."""
    import numpy as np
    out, parity = {}, {}

    def best(fn, reps=2):
        b = float("inf")
        for _ in range(reps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            r = fn()
            torch.cuda.synchronize()
            b = min(b, time.perf_counter() - t0)
        return b, r

    def pipeline(propagate, pts):
        propagate(pts)
        if bool(field_utils.measure_mean_potential(pts) < 0):
            pts[:, 3:] *= -1
        return pts

    def flips_of(pts, pc_in):
        return np.packbits(((pts[:, 3:] * pc_in[:, 3:]).sum(-1) < 0).numpy())

    g = _golden("points_prop")
    pc = torch.from_numpy(g["pc_in"].copy())
    for key, n, diffuse, tag in (("cfg1_ok_xyz_first_2000_pts_e2e_s", 2000, True, "n2000"),
                                 ("cfg1_ok_xyz_10000_pts_e2e_s", 10000, False, "full")):
        sec, res = best(lambda: pipeline(lambda p: field_utils.strongest_field_propagation_points(p, diffuse=diffuse),
                                         pc[:n].clone()))
        out[key] = sec
        want = torch.from_numpy(g[f"{tag}_normals"].copy())
        if float(g[f"{tag}_mean_potential"]) < 0:
            want = -want
        parity[key] = bool(torch.equal(res[:, 3:], want))
    g2 = _golden("patch_prop")
    groups = _csr_patches(g2["perm"], g2["offsets"])
    pc2 = torch.from_numpy(g2["pc_in"].copy())
    sec, res = best(lambda: pipeline(lambda p: field_utils.strongest_field_propagation(p, [], groups), pc2.clone()))
    out["cfg2_fandisk_72_patches_e2e_s"] = sec
    want = torch.from_numpy(g2["plain_normals"].copy())
    if float(g2["plain_mean_potential"]) < 0:
        want = -want
    parity["cfg2_fandisk_72_patches_e2e_s"] = bool(torch.equal(res[:, 3:], want))
    try:
        gl = _golden("large")
    except OSError:
        gl = None
    pc3, reps = cfg3_standin()
    sec, res = best(lambda: pipeline(lambda p: field_utils.strongest_field_propagation_reps(p, reps, diffuse=True),
                                     pc3.clone()))
    out["cfg3_reps_60000_pts_48_patches_e2e_s"] = sec
    if gl is not None and abs(pc3.double().sum().item() - float(gl["cfg3_input_checksum"])) < 1e-9:
        got = flips_of(res, pc3)
        want = gl["cfg3_diffuse_flip_bits"]
        if float(gl["cfg3_diffuse_mean_potential"]) < 0:
            want = np.packbits(1 - np.unpackbits(want)[:pc3.shape[0]])
        parity["cfg3_reps_60000_pts_48_patches_e2e_s"] = {
            "points_differing_from_reference": int(np.unpackbits(got ^ want)[:pc3.shape[0]].sum()), "points": pc3.shape[0]}
    ref, tgt = cfg4_standin()
    sec, res = best(lambda: field_utils.reference_field(ref, tgt.clone()))
    out["cfg4_transfer_20000_x_40000_e2e_s"] = sec
    if gl is not None and abs(ref.double().sum().item() + tgt.double().sum().item() - float(gl["cfg4_input_checksum"])) < 1e-9:
        diff = np.unpackbits(flips_of(res, tgt) ^ gl["cfg4_flip_bits"])[:tgt.shape[0]].astype(bool)
        parity["cfg4_transfer_20000_x_40000_e2e_s"] = {
            "points_differing_from_reference": int(diff.sum()), "points": tgt.shape[0],
            "largest_margin_among_them": float(gl["cfg4_margin"][diff].max()) if diff.any() else 0.0}
    out["parity_vs_reference_decisions"] = parity
    return out


def run_reference_arm(args, rank, world):
    """`--impl reference`: the reference's CPU torch path (ported op for op, see oracle/torch_port.py) on this
    box's host cores.  Rank 0 alone runs; each step is a bounded sample of the workload."""
    if rank != 0:
        return
    torch.set_num_threads(os.cpu_count() or 1)
    n = args.cpu_points
    for _ in range(args.warmup):
        time_cpu_port(n, 2 * n, 1)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        time_cpu_port(n, 2 * n, 1)
    sec = time.perf_counter() - t0
    value = args.steps * n * 2 * n / sec
    sample = (f"each step = oracle/torch_port.field_grad on {n} x {2 * n} float32 points ({2 * n * n:.3e} pairs, the "
              f"targets above max_pts = 15000 so that the reference's recursive chunking runs), a bounded sample of "
              f"the {args.points} x {args.points} per-GPU workload; the reference is pure Python/torch and does not "
              f"travel to this box, so its op chain is ported (bit-exact vs tests/golden/leaf.npz)")
    extras = {} if args.no_extras else cpu_driver_seconds(full=True)
    if not args.no_extras:       # field_utils.potential (field_utils.py:12-55) on the same sample shape, once
        extras["potential_pairs_per_s"] = n * 2 * n / time_cpu_port(n, 2 * n, potential=True)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, world),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                         "host_cpus": os.cpu_count(), "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "extras": extras, "gpu_launches": 0}), flush=True)


def ncu_traffic(points):
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture (profiles/),
    or None when no capture exists for this size."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[f"upairs_tile_kernel_field_896x{points}"]
        return t["dram_bytes_read"] + t["dram_bytes_write"]
    except (OSError, KeyError, ValueError):
        return None


def workload_config(args, world):
    return {"workload": f"field_grad all-pairs sweep (BASELINE config 5): {args.points} replicated sources x "
                        f"{args.points} targets per GPU, eps={EPS}, uniform synthetic cloud",
            "sources": args.points, "targets_per_gpu": args.points, "targets_total": args.points * world,
            "sharding": "targets block-sharded over ranks, sources replicated, no data-path collective",
            "l2": "256 MiB scratch buffer rewritten between timed steps (inputs are smaller than the 126 MB L2; "
                  "the kernel is FP32-pipe bound, not memory bound)"}


# ---------------------------------------------------------------------------------------------- GPU arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--points", type=int, default=1 << 20, help="sources (replicated) and targets per GPU")
    ap.add_argument("--cpu-points", type=int, default=8192, help="side of the CPU sample (reference arm / baseline)")
    ap.add_argument("--strong-points", type=int, default=1 << 21,
                    help="side of the fixed global problem of the strong-scaling block (dnp_dist)")
    ap.add_argument("--no-strong", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    import torch.distributed as dist
    import dnp_native as dn
    import field_utils

    dn.require_cuda()                                   # no CPU fallback: fail loudly without a GPU
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    parity = None
    if world > 1:
        import dnp_dist
        parity = dist_parity(dn, dist, dnp_dist, dev, rank, world)

    P = args.points
    src_host = cloud(P, 1, pin=True)                                    # replicated sources
    tgt_host = cloud(P, 2000 + rank, cols=3, pin=True)                 # this rank's target shard
    src, tgt = src_host.to(dev), tgt_host.to(dev)
    E = torch.empty(P, 3, dtype=torch.float32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    pairs_per_step = float(P) * float(P)

    def step():
        flush.zero_()
        dn.field_grad(src, tgt, eps=EPS, out=E)

    # roofline denominator measured on this box in this process: FP32 FMA issue rate (packed FFMA2)
    mb_ms, mb_ops = dn.microbench(1, 4000)
    fma_per_s = mb_ops / (mb_ms * 1e-3)
    sfu_ms, sfu_ops = dn.microbench(2, 4000)
    info = dn.device_info()

    for _ in range(args.warmup):
        step()
    barrier()
    launches0 = dn.launch_count()
    dn.profile_begin()
    with ClockSampler(local_rank) as clocks:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(args.steps):
            step()
        e1.record()
        barrier()
    kernel_ms = dn.profile_end()
    launches = dn.launch_count() - launches0
    ms_total = e0.elapsed_time(e1)

    # ---- end to end through the reference-facing API with host buffers
    def e2e_step():
        return field_utils.field_grad(src_host, tgt_host, eps=EPS)

    e2e_step()
    barrier()
    with ClockSampler(local_rank) as e2e_clocks:
        t0 = time.perf_counter()
        e2e_calls = []
        for _ in range(args.steps):
            t1 = time.perf_counter()
            out_host = e2e_step()
            e2e_calls.append((time.perf_counter() - t1) * 1e3)
        barrier()
        e2e_s = time.perf_counter() - t0
    if os.environ.get("DNP_BENCH_DEBUG"):
        print("e2e per call ms:", ["%.1f" % v for v in e2e_calls], file=sys.stderr)
    assert not out_host.is_cuda and out_host.shape == (P, 3)

    t = torch.tensor([ms_total, e2e_s * 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, e2e_ms = t.tolist()

    strong = None
    if not args.no_strong:
        del E, flush
        strong = multi_gpu_blocks(args, dn, dist, dev, rank, world)

    if rank == 0:
        value = world * pairs_per_step * args.steps / (ms_total * 1e-3)
        # one timed region per step: the fork->join span of the step's chunk launches (dnp_upairs.cu runs them on
        # overlapping lanes, so single-launch durations are not separable; the region is what the events see)
        k_ms = sum(kernel_ms) / max(len(kernel_ms), 1)
        chunk_launches = (P + 895) // 896                             # window slots of 896 sources (dnp_upairs.cu)
        achieved = pairs_per_step * FLOP_PER_PAIR_FIELD / (k_ms * 1e-3) / 1e12
        peak = 2.0 * fma_per_s / 1e12
        nominal = info["sm_count"] * 128 * 2 * info["sm_clock_khz"] * 1e3 / 1e12
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, world),
            "e2e": {"value": world * pairs_per_step * args.steps / (e2e_ms * 1e-3), "unit": UNIT,
                    "ms_per_call": [round(v, 1) for v in e2e_calls], "clocks": e2e_clocks.summary(),
                    "h2d_bytes_per_step": int(src_host.numel() * 4 + tgt_host.numel() * 4),
                    "d2h_bytes_per_step": int(P * 3 * 4),
                    "api": "field_utils.field_grad(host float32 tensors) -> dnp_field_grad_host_f32"},
            "gpu_launches": int(launches),
            "clocks": clocks.summary(),
            "roofline": {
                "bound": "fp32", "kernel": "dnp::upairs_tile_kernel<false> (all-pairs field, sources in uniform registers "
                                           "fed from the device-filled constant window; one launch = 896 sources x all "
                                           "targets)",
                "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                "peak_source": "measured on this GPU in this process: packed FFMA2 issue microbenchmark "
                               "(dnp_microbench kind 1) x 2 flop; MEASURED_PEAKS.json holds only HBM and bf16 peaks",
                "nominal_peak": nominal, "frac_of_nominal": achieved / nominal,
                "flop_per_pair": FLOP_PER_PAIR_FIELD, "pairs_per_launch": 896.0 * P,
                "launches_per_step": chunk_launches, "kernel_ms_avg": k_ms / chunk_launches,
                "region_ms_avg": k_ms, "regions_timed": len(kernel_ms),
                "kernel_share_of_step": k_ms * len(kernel_ms) / ms_total if ms_total else None,
                "pairs_per_s_kernel": pairs_per_step / (k_ms * 1e-3),
                "fp32_instr_per_pair_executed": 19, "sfu_per_pair_executed": 2,
                "sfu_peak_gops": sfu_ops / (sfu_ms * 1e-3) / 1e9,
                "sfu_frac": 2.0 * pairs_per_step / (k_ms * 1e-3) / (sfu_ops / (sfu_ms * 1e-3)),
                "hbm_algorithmic_gbs": chunk_launches * (P * 12 + 2 * P * 12) / (k_ms * 1e-3) / 1e9,
                "traffic": ncu_traffic(P),
                "traffic_note": "DRAM bytes of ONE launch under ncu, which flushes the caches before it; all 1171 launches "
                                "of a step move 82 MB in total with warm caches (profiles/r02_dram_warm_summary.txt) "
                                "against 50 MB algorithmic"},
        }
        if parity is not None:
            line["dist_parity"] = parity
        if strong is not None:
            one_gpu_rate = value / world                     # this run's own per-GPU rate on the weak workload
            f = strong["field"]
            f["speedup_vs_one_gpu_rate"] = f["pairs_per_s"] / one_gpu_rate
            f["efficiency_vs_one_gpu_rate"] = f["pairs_per_s"] / one_gpu_rate / world
            line["strong"] = dict(strong, scaling="strong", n_gpus=world)
        if not args.no_extras:
            line["extras"] = extras(dn, field_utils, dev, src, tgt)
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_block(args.cpu_points)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def extras(dn, field_utils, dev, src, tgt):
    """Other rows of the hot path, timed outside the headline region (CUDA events, best of 3)."""
    import numpy as np

    def best_ms(fn, reps=3):
        fn()
        torch.cuda.synchronize()
        best = float("inf")
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b))
        return best

    P = src.shape[0]
    out = gpu_driver_seconds(field_utils, dev)
    phi = torch.empty(P, device=dev)
    ms = best_ms(lambda: dn.potential(src, tgt, out=phi), 2)
    out["potential_pairs_per_s"] = P * P / ms * 1e3
    out["potential_frac_fp32_nominal"] = P * P / ms * 1e3 * FLOP_PER_PAIR_POTENTIAL / (148 * 128 * 2 * 1.965e9)
    ms = best_ms(lambda: dn.lattice_mean_potential(src))
    out["lattice_test_ms"] = ms
    out["cfg5_sweep"] = cfg5_sweep(dn, dev)
    out["lattice_pairs_per_s"] = P * 1000 / ms * 1e3
    # streamed sign phase (HBM-bound): 24 B read + 12 B written per point
    big = torch.randn(1 << 24, 6, device=dev)
    Eb = torch.randn(1 << 24, 3, device=dev)
    ms = best_ms(lambda: dn.field_sign(Eb, big))
    out["field_sign_gbs"] = (1 << 24) * 36 / ms / 1e6
    del big, Eb
    # cfg1 / cfg2 drivers on the shipped demo clouds (inputs from the committed golden fixtures)
    gdir = os.path.join(ROOT, "tests", "golden")
    try:
        g = np.load(os.path.join(gdir, "points_prop.npz"))
        pc = torch.from_numpy(g["pc_in"]).to(dev)

        def run_points():
            work = pc.clone()
            field_utils.strongest_field_propagation_points(work)
            return field_utils.measure_mean_potential(work)
        out["cfg1_ok_xyz_points_propagation_ms"] = best_ms(run_points)
        out["cfg1_reference_cpu_s_build_container"] = float(g["full_ref_seconds"])
        g = np.load(os.path.join(gdir, "patch_prop.npz"))
        pc2 = torch.from_numpy(g["pc_in"]).to(dev)
        patches = [torch.from_numpy(g["perm"][g["offsets"][i]:g["offsets"][i + 1]]).to(dev)
                   for i in range(int(g["P"]))]

        def run_patches():
            work = pc2.clone()
            field_utils.strongest_field_propagation(work, [], patches)
            return field_utils.measure_mean_potential(work)
        out["cfg2_fandisk_patch_propagation_ms"] = best_ms(run_patches)
        out["cfg2_reference_cpu_s_build_container"] = float(g["plain_ref_seconds"])
    except OSError:
        pass
    # cfg1 at the size of the shipped 100 000-point clouds (flower/galera/boxunion.xyz): cooperative-grid kernel
    big_pc = surface_cloud(100000, 7).to(dev)
    out["points_propagation_100k_ms"] = best_ms(lambda: field_utils.strongest_field_propagation_points(big_pc.clone()), 1)
    # cfg4 stand-in (data/interpolate/*.xyz are not shipped): 100 000 oriented reference points -> 1M-point target
    ref_pc = surface_cloud(100000, 8).to(dev)
    tgt3 = surface_cloud(1 << 20, 9)[:, :3].contiguous().to(dev)
    out["cfg4_reference_field_100k_x_1M_ms"] = best_ms(lambda: field_utils.reference_field(ref_pc, tgt3), 2)
    # cfg3 stand-in (data/lion.xyz is not shipped): 1M surface points, 1000 spatially coherent patches of ~1000
    # points, 500 representatives each (orient_large.py:48-52), diffuse, then the lattice test
    pc3 = surface_cloud(1 << 20, 10)
    key = ((pc3[:, :3] + 0.6) * 8).long().clamp(0, 15)
    order = torch.argsort(key[:, 0] * 256 + key[:, 1] * 16 + key[:, 2], stable=True)
    chunks = torch.chunk(order, 1000)
    reps = [(c[:500].to(dev), c[500:].to(dev)) for c in chunks]
    pc3 = pc3.to(dev)

    def run_reps():
        work = pc3.clone()
        field_utils.strongest_field_propagation_reps(work, reps, diffuse=True)
        return field_utils.measure_mean_potential(work)
    out["cfg3_reps_propagation_1M_pts_1000_patches_ms"] = best_ms(run_reps, 1)
    # xie in-order propagation, 5 chains over 10 000 points (the server's tree_xie call)
    pcx = surface_cloud(10000, 11).to(dev)
    gx = torch.Generator().manual_seed(12)
    orders = torch.stack([torch.randperm(10000, generator=gx) for _ in range(5)]).to(dev)
    out["xie_in_order_10k_pts_5_chains_diffuse_ms"] = best_ms(
        lambda: field_utils.xie_propagation_points_in_order(pcx, 1e-2, orders, diffuse=True), 2)
    ms = best_ms(lambda: dn.xie_matrix(pcx, pcx), 2)
    out["xie_matrix_10k_gbs_written"] = 10000 * 10000 * 4 / ms / 1e6
    # voxel partition + small-patch merge of the 1M-point cloud at --number_parts 41
    import dnp_util
    ms = best_ms(lambda: dn.voxel_partition(pc3, 41), 3)
    out["partition_1M_pts_41_parts_ms"] = ms
    out["partition_algorithmic_gbs"] = (1 << 20) * 20 / ms / 1e6        # 12 B/point read + 8 B/point of perm written
    dnp_util.divide_pc(pc3, 41, min_patch=100)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    merged = dnp_util.divide_pc(pc3, 41, min_patch=100)
    torch.cuda.synchronize()
    out["partition_plus_merge_python_api_ms"] = (time.perf_counter() - t0) * 1e3
    out["partition_patches"] = len(merged)
    out["partition_reference_cpu_s_build_container"] = "5.8 s for 10 000 points at n_part 41 (tests/golden/partition.npz, ok_divide_seconds)"
    # patch-graph flip solve (graph.MIQP): exhaustive walk at 36 patches, annealing at 300
    gq = torch.Generator().manual_seed(99)

    def couplings(n, density):
        w = torch.triu(torch.randn(n, n, generator=gq, dtype=torch.float64), 1)
        w = w * (torch.rand(n, n, generator=gq) < density)
        return (w + w.T).to(dev)
    a36, a300 = couplings(36, 0.4), couplings(300, 0.03)
    ms = best_ms(lambda: dn.flip_solve(a36, -a36), 2)
    out["flip_solve_36_patches_exhaustive_ms"] = ms
    out["flip_solve_exhaustive_states_per_s"] = 2.0 ** 35 / (ms * 1e-3)
    out["flip_solve_300_patches_anneal_ms"] = best_ms(lambda: dn.flip_solve(a300, -a300), 2)
    # the patch graph of a 100k-point cloud (grid 10, as graph_dipole.py:131-135):
    # per-patch greedy orientation (one launch, a cluster per patch, vs one call per patch) + batched flip scores
    import contextlib
    import io
    import graph_dipole
    pcg = surface_cloud(100000, 12).to(dev)
    pcg[:, :3] += 0.013                                             # off the cell faces

    def run_graph(estimator):
        with contextlib.redirect_stdout(io.StringIO()):
            return dnp_util.divide_pc_to_graph(pcg.clone(), n_part=10, min_patch=50,
                                               edge_calculator=field_utils.field_edge_calculator,
                                               point_estimator=estimator)
    G, index = run_graph(graph_dipole.st_propagation_points_file)
    out["patch_graph_100k_pts_patches"] = len(index)
    out["patch_graph_100k_pts_edges"] = len(G.E)
    t0 = time.perf_counter(); run_graph(graph_dipole.st_propagation_points_file); torch.cuda.synchronize()
    out["patch_graph_100k_pts_batched_ms"] = (time.perf_counter() - t0) * 1e3
    t0 = time.perf_counter(); run_graph(lambda p: graph_dipole.st_propagation_points_file(p, verbose=False)); torch.cuda.synchronize()
    out["patch_graph_100k_pts_per_patch_calls_ms"] = (time.perf_counter() - t0) * 1e3
    # text IO of the 1M-point cloud (export -> parse), device time and end to end from host bytes
    cloud = pc3[:, :6] if pc3.shape[1] >= 6 else torch.cat([pc3, torch.zeros_like(pc3)], 1)
    rows = cloud.contiguous()
    import tempfile
    dnp_util.export_text(rows.transpose(0, 1), as_view=True)
    t0 = time.perf_counter()
    view = dnp_util.export_text(rows.transpose(0, 1), as_view=True)                  # device rows -> text in host memory
    out["xyz_export_1M_pts_e2e_ms"] = (time.perf_counter() - t0) * 1e3
    del view
    with tempfile.TemporaryDirectory() as tmp:
        times = []
        for k in range(3):                                       # a new file every time; the first call also pins 8 MB
            t0 = time.perf_counter()
            dnp_util.export_pc(rows.transpose(0, 1), os.path.join(tmp, f"cloud{k}.xyz"))   # util.export_pc: ... into a file
            times.append((time.perf_counter() - t0) * 1e3)
        out["xyz_export_1M_pts_to_file_ms"] = min(times)
        out["xyz_export_1M_pts_to_file_first_call_ms"] = times[0]
    text = dnp_util.export_text(rows.transpose(0, 1))
    out["xyz_text_bytes"] = len(text)
    out["xyz_format_1M_pts_ms"] = best_ms(lambda: _format_only(dn, rows), 2)
    t0 = time.perf_counter()
    back, mask, bad, lines = dn.xyz_parse(text)
    torch.cuda.synchronize()
    out["xyz_parse_1M_pts_e2e_ms"] = (time.perf_counter() - t0) * 1e3
    txt_dev = torch.frombuffer(bytearray(text), dtype=torch.uint8).to(dev)
    out["xyz_parse_1M_pts_ms"] = best_ms(lambda: _parse_only(dn, txt_dev, lines), 2)
    out["xyz_parse_text_gbs"] = len(text) / out["xyz_parse_1M_pts_ms"] / 1e6
    out["xyz_round_trip_bit_exact"] = bool(torch.equal(back.view(torch.int32), rows.view(torch.int32)))
    out["xyz_reference_cpu"] = "util.xyz2tensor / export_pc are per-line Python loops: 1.2 us per value parsed, 1.5 us per value written on the build container (tests/test_io.py::test_reference_style_loops_rate), i.e. 7 s + 9 s for 1M points"
    return out


def _format_only(dn, rows):
    lib = dn.load()
    n, c = rows.shape
    cap = 25 * n * c
    dev = rows.device
    out = torch.empty(cap, dtype=torch.uint8, device=dev)
    n_out = torch.zeros(1, dtype=torch.int64, device=dev)
    nbytes = lib.dnp_xyz_format_workspace_bytes(n * c)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    dn._check(lib.dnp_xyz_format_f32(dn._ptr(rows), n, c, rows.stride(0), rows.stride(1), dn._ptr(out), cap,
                                     dn._ptr(n_out), dn._ptr(ws), nbytes, dn._stream(rows)))


def _parse_only(dn, txt_dev, n_lines):
    lib = dn.load()
    dev = txt_dev.device
    out = torch.empty(n_lines, 6, dtype=torch.float32, device=dev)
    counts = torch.empty(4, dtype=torch.int64, device=dev)
    nbytes = lib.dnp_xyz_parse_workspace_bytes(txt_dev.numel(), n_lines)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    dn._check(lib.dnp_xyz_parse_f32(dn._ptr(txt_dev), txt_dev.numel(), n_lines, dn._ptr(out), n_lines, dn._ptr(counts),
                                    dn._ptr(ws), nbytes, dn._stream(txt_dev)))


def cfg5_sweep(dn, dev):
    """BASELINE config 5 / SURVEY.md section 8(d): the synthetic all-pairs sweep at its other sizes (1M x 1M is the
    headline), each with the rate, the fraction of the nominal FP32 peak and the error of sampled targets against the
    same sum evaluated in float64 (torch on the device: a referee for the arithmetic, not the oracle)."""
    eps = 1e-5
    rows = []
    for k, (S, M) in enumerate([(65536, 65536), (262144, 262144), (65536, 4194304), (4194304, 4194304)]):
        g = torch.Generator().manual_seed(1234 + k)
        src = torch.rand(S, 6, generator=g) - 0.5
        src[:, 3:] = torch.nn.functional.normalize(src[:, 3:], dim=1)
        tgt = torch.rand(M, 3, generator=g) - 0.5
        src, tgt = src.to(dev), tgt.to(dev)
        row = {"sources": S, "targets": M}
        for name, flop in (("field", FLOP_PER_PAIR_FIELD), ("potential", FLOP_PER_PAIR_POTENTIAL)):
            out = torch.empty((M, 3) if name == "field" else (M,), device=dev)
            call = (lambda: dn.field_grad(src, tgt, eps=eps, out=out)) if name == "field" else \
                (lambda: dn.potential(src, tgt, out=out))
            big = S * M > 4e12
            if not big:
                call()
            best = float("inf")
            for _ in range(1 if big else 3):
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                call()
                b.record()
                torch.cuda.synchronize()
                best = min(best, a.elapsed_time(b))
            rate = S * M / best * 1e3
            row[name] = {"ms": best, "pairs_per_s": rate, "frac_fp32_nominal": rate * flop / (148 * 128 * 2 * 1.965e9),
                         "timed_runs": 1 if big else 3}
            # float64 referee on 64 evenly spaced targets
            pick = torch.linspace(0, M - 1, 64, device=dev).long()
            q = tgt[pick].double()
            ref = torch.zeros((64, 3) if name == "field" else (64,), device=dev, dtype=torch.float64)
            mag = torch.zeros(64, device=dev, dtype=torch.float64)
            for c0 in range(0, S, 1 << 19):
                sc = src[c0:c0 + (1 << 19)].double()
                R = sc[:, None, :3] - q[None]
                r = R.norm(dim=-1)
                n = sc[:, None, 3:]
                if name == "field":
                    Ru = torch.where(r[..., None] > 0, R / r[..., None].clamp_min(1e-300), torch.zeros_like(R))
                    e = (3.0 * (n * Ru).sum(-1, keepdim=True) * Ru - n) / (r ** 3 + eps)[..., None]
                    e = torch.where(r[..., None] > 0, e, torch.zeros_like(e))
                    ref -= e.sum(0)
                else:
                    ph = torch.where(r > 0, (n * R).sum(-1) / r.clamp_min(1e-300) ** 3, torch.zeros_like(r))
                    ref += ph.sum(0)
                    mag += ph.abs().sum(0)
            got = out[pick].double()
            if name == "field":
                err = (got - ref).norm(dim=1) / ref.norm(dim=1).clamp_min(1e-300)
            else:
                # the potential of a random cloud passes through 0: scale by the sum of the magnitudes
                err = (got - ref).abs() / mag.clamp_min(1e-300)
            row[name]["rel_err_vs_f64_median"] = float(err.median())
            row[name]["rel_err_vs_f64_max"] = float(err.max())
            row[name]["rel_err_scale"] = "|E64| per target" if name == "field" else "sum of |phi_s| per target"
            del out
        rows.append(row)
        del src, tgt
    return rows


def surface_cloud(n, seed):
    """Noisy unit-normal sphere of radius 0.45 with a seeded 50 % sign scramble."""
    g = torch.Generator().manual_seed(seed)
    d = torch.randn(n, 3, generator=g)
    d = d / d.norm(dim=-1, keepdim=True)
    pc = torch.cat([d * 0.45 + torch.randn(n, 3, generator=g) * 0.002, d], 1).float()
    pc[torch.rand(n, generator=g) < 0.5, 3:] *= -1
    return pc.contiguous()


def cfg3_standin(n=60000, n_patches=48, max_rep=500, seed=10):
    """BASELINE config 3 stand-in at the size BASELINE.md section 2 timed the reference (data/lion.xyz is not shipped):
    a scrambled noisy sphere, `n_patches` spatially coherent patches (points sorted by a coarse voxel key, equal
    chunks), `max_rep` seeded random representatives per patch (orient_large.py:48-52).  CPU tensors."""
    pc = surface_cloud(n, seed)
    key = ((pc[:, :3] + 0.6) * 8).long().clamp(0, 15)
    order = torch.argsort(key[:, 0] * 256 + key[:, 1] * 16 + key[:, 2], stable=True)
    g = torch.Generator().manual_seed(seed + 1)
    reps = []
    for c in torch.chunk(order, n_patches):
        pm = c[torch.randperm(c.shape[0], generator=g)]
        reps.append((pm[:max_rep].clone(), pm[max_rep:].clone()))
    return pc, reps


def cfg4_standin(n_ref=20000, n_tgt=40000, seed=8):
    """BASELINE config 4 stand-in (data/interpolate/*.xyz are not shipped), the size BASELINE.md section 2 timed:
    an oriented reference sampling of the sphere and a scrambled, independently sampled target.  CPU tensors."""
    g = torch.Generator().manual_seed(seed)
    d = torch.randn(n_ref, 3, generator=g)
    d = d / d.norm(dim=-1, keepdim=True)
    ref = torch.cat([d * 0.45 + torch.randn(n_ref, 3, generator=g) * 0.002, d], 1).float().contiguous()
    return ref, surface_cloud(n_tgt, seed + 1)


if __name__ == "__main__":
    main()
