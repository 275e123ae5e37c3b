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
def time_cpu_port(n_src, n_tgt, repeats=1):
    """Seconds per field_grad(n_src x n_tgt) of the torch port on all host cores (best of `repeats`)."""
    from oracle import torch_port
    src, tgt = cloud(n_src, 1), cloud(n_tgt, 2, cols=3)
    best = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        torch_port.field_grad(src, tgt, eps=EPS)
        best = min(best, time.perf_counter() - t0)
    return best


def cpu_baseline_block(n):
    torch.set_num_threads(os.cpu_count() or 1)
    time_cpu_port(2048, 2048)                          # warm the thread pool and the allocator
    m = 2 * n                                          # > max_pts = 15000 at the default n: the chunking runs too
    sec = time_cpu_port(n, m)
    return {"value": n * m / sec, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
            "host_cpus": os.cpu_count(),
            "sample": f"oracle/torch_port.field_grad (the reference's op chain and chunking, field_utils.py:61-116), "
                      f"float32, {n} sources x {m} targets = {n * m:.3e} pairs, one run of {sec:.1f} s"}


def run_reference_arm(args, rank, world):
    """`--impl reference`: the reference's CPU torch path (ported op for op, see oracle/torch_port.py) on this
    box's host cores.  Rank 0 alone runs; each step is a bounded sample of the workload."""
    if rank != 0:
        return
    torch.set_num_threads(os.cpu_count() or 1)
    n = args.cpu_points
    for _ in range(args.warmup):
        time_cpu_port(n, n, 1)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        time_cpu_port(n, n, 1)
    sec = time.perf_counter() - t0
    value = args.steps * n * n / sec
    sample = (f"each step = oracle/torch_port.field_grad on {n} x {n} float32 points ({n * n:.3e} pairs), a bounded "
              f"sample of the {args.points} x {args.points} per-GPU workload; the reference is pure Python/torch and "
              f"does not travel to this box, so its op chain is ported (bit-exact vs tests/golden/leaf.npz)")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, world),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                         "host_cpus": os.cpu_count(), "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}), flush=True)


def ncu_traffic(points):
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture (profiles/),
    or None when no capture exists for this size."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[f"cpairs_kernel_field_896x{points}"]
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

    if rank == 0:
        value = world * pairs_per_step * args.steps / (ms_total * 1e-3)
        # one timed region per step: the fork->join span of the step's chunk launches (dnp_cpairs.cu runs them on
        # overlapping lanes, so single-launch durations are not separable; the region is what the events see)
        k_ms = sum(kernel_ms) / max(len(kernel_ms), 1)
        chunk_launches = ((P + 223) // 224 * 224 + 895) // 896        # blocks of 896 sources (dnp_cpairs.cu)
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
                "bound": "fp32", "kernel": "dnp::cpairs_kernel<false> (all-pairs field, sources by value in uniform "
                                           "registers; one launch = 896 sources x all targets)",
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
                "traffic": ncu_traffic(P)},
        }
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
    out = {}
    phi = torch.empty(P, device=dev)
    ms = best_ms(lambda: dn.potential(src, tgt, out=phi), 2)
    out["potential_pairs_per_s"] = P * P / ms * 1e3
    out["potential_frac_fp32_nominal"] = P * P / ms * 1e3 * FLOP_PER_PAIR_POTENTIAL / (148 * 128 * 2 * 1.965e9)
    ms = best_ms(lambda: dn.lattice_mean_potential(src))
    out["lattice_test_ms"] = ms
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
    t0 = time.perf_counter()
    idx, cells = dnp_util._divide_pc(pc3[:, :3], 41)
    merged, _ = dnp_util.merge_nodes(pc3, idx, cells, 100)
    torch.cuda.synchronize()
    out["partition_plus_merge_python_api_ms"] = (time.perf_counter() - t0) * 1e3
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
    t0 = time.perf_counter()
    text = dnp_util.export_text(rows.transpose(0, 1))
    out["xyz_export_1M_pts_e2e_ms"] = (time.perf_counter() - t0) * 1e3
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


def surface_cloud(n, seed):
    """Noisy unit-normal sphere of radius 0.45 with a seeded 50 % sign scramble."""
    g = torch.Generator().manual_seed(seed)
    d = torch.randn(n, 3, generator=g)
    d = d / d.norm(dim=-1, keepdim=True)
    pc = torch.cat([d * 0.45 + torch.randn(n, 3, generator=g) * 0.002, d], 1).float()
    pc[torch.rand(n, generator=g) < 0.5, 3:] *= -1
    return pc.contiguous()


if __name__ == "__main__":
    main()
