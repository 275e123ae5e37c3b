"""Torch (CPU) restatement of the reference's four DRIVERS on the hot path, with the reference's own ATen op
sequence.  TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): it is what `bench.py` times on the GPU box's host
cores as the CPU side of "e2e orient sec" (`cpu_baseline.drivers`, `--impl reference`), and a checker in tests/.

The reference is pure Python over torch and does not travel to the GPU box; these functions keep its op chain so
that (a) the float32 results are the reference's bit for bit (pinned by tests/test_oracle.py against
tests/golden/{points_prop,patch_prop,reps_prop,transfer}.npz, which the unmodified reference generated) and
(b) the time they take is the time the reference takes: the cost of the reference's drivers is the Python loop
with its boolean-mask gathers, per-patch list comprehension and host synchronisations, not the flops.

    transfer_orientation      <- field_utils.reference_field                      field_utils.py:188-201
    propagate_points          <- field_utils.strongest_field_propagation_points   field_utils.py:353-388
    propagate_patches         <- field_utils.strongest_field_propagation          field_utils.py:286-348
    propagate_representatives <- field_utils.strongest_field_propagation_reps     field_utils.py:207-282
    lattice_mean_potential    <- field_utils.measure_mean_potential + util.gen_grid   field_utils.py:7-9, util.py:26-36
    smallest_pca_eigenvalue   <- util.pca_eigen_values                            util.py:495-500

Each function takes an optional `trace` list that receives the visit order (patch / point ids), which the
reference does not return (tests/golden/make_golden.py recovers it by wrapping field_grad)."""
import numpy as np
import torch

from .torch_port import field_grad, potential


def lattice_nodes(n=10):
    """util.gen_grid (util.py:26-36): n^3 nodes, x-major, coordinates (i/n - 0.5) * 2."""
    k = torch.arange(n ** 3)
    z = k % n
    xy = k // n
    y = xy % n
    x = xy // n
    return (torch.stack([x, y, z], dim=1).float() / n - 0.5) * 2


def lattice_mean_potential(pc):
    """field_utils.py:7-9."""
    return potential(pc, lattice_nodes().to(pc.device)).mean()


def smallest_pca_eigenvalue(x):
    """util.py:495-500 (torch.symeig -> linalg.eigh, same ascending order): lambda_0 of the xyz covariance."""
    centred = x[:, :3] - x.mean(dim=0)[None, :3]
    cov = (centred.transpose(0, 1) @ centred) / x.shape[0]
    lam, _ = torch.linalg.eigh(cov)
    return lam[0:1]


def _flattest(pts, groups):
    """field_utils.py:230-233 / :303-306: argmin |lambda_0| over the patches, first on ties."""
    lam0 = np.array([smallest_pca_eigenvalue(pts[g])[0].cpu() for g in groups])
    return int(np.argmin(np.abs(lam0)))


def _sign_of(dots):
    return (dots > 0).float() * 2 - 1


def transfer_orientation(oriented, target):
    """field_utils.py:188-201.  3-column target: append E/|E|; 6-column: flip in place where E.n < 0."""
    with torch.no_grad():
        E = field_grad(oriented, target, recursive=True)
        if target.shape[1] == 3:
            mag = E.norm(dim=-1)
            E[mag != 0, :] = E[mag != 0, :] / mag[mag != 0, None]
            return torch.cat([target, E], dim=1)
        dots = (E * target[:, 3:]).sum(dim=-1)
        keep = (dots >= 0).float() * 2 - 1
        target[:, 3:] = target[:, 3:] * keep[:, None]
        return target


def propagate_points(pts, diffuse=False, starting_point=0, trace=None, flips=None, max_steps=None):
    """field_utils.py:353-388 (the `.cuda()` of :355 is the identity on the CPU path).  N - 1 steps, each a
    1 x (N - 1) field_grad with eps 1e-6, a masked sum, abs().argmax() and a conditional flip.
    max_steps: stop after that many greedy steps (bounded timing sample)."""
    ids = torch.arange(pts.shape[0])
    E = torch.zeros_like(pts[:, :3])
    done = torch.zeros_like(pts[:, 0]).bool()
    done[starting_point] = True
    others = ~(ids == starting_point)
    E[others] += field_grad(pts[starting_point:(starting_point + 1)], pts[others, :3], eps=1e-6)
    if trace is not None:
        trace.append(int(starting_point))
    steps = 0
    while not done.all():
        if max_steps is not None and steps >= max_steps:
            break
        score = (E[~done] * pts[~done, 3:]).sum(dim=-1)
        k = score.abs().argmax()
        who = ids[~done][k]
        flip = bool(score[k] < 0)
        if flip:
            pts[who, 3:] *= -1
        done[who] = True
        if trace is not None:
            trace.append(int(who))
        if flips is not None:
            flips.append(flip)
        others = ~(ids == who)
        E[others] += field_grad(pts[who:(who + 1)], pts[others, :3], eps=1e-6)
        steps += 1
    if diffuse:
        pts[:, 3:] = pts[:, 3:] * _sign_of((E * pts[:, 3:]).sum(dim=-1))[:, None]
    return pts


def propagate_patches(pts, kept, groups, diffuse=False, weights=None, trace=None):
    """field_utils.py:286-348.  `groups` = every patch (index tensors), `kept` = [(i, idx)] the patches whose
    points take the final per-point sign when diffuse."""
    with torch.no_grad():
        if weights is not None:
            weights = weights.clamp(0.1, 1)
            pts[:, 3:] = pts[:, 3:] * weights[:, None]
        todo = [(i, groups[i]) for i in range(len(groups))]
        oriented = torch.zeros(len(pts)).bool()
        nothing = torch.zeros(len(pts)).bool()
        E = torch.zeros_like(pts[:, :3])
        first, rows = todo.pop(_flattest(pts, groups))
        if trace is not None:
            trace.append(first)
        oriented[rows] = True
        E[~oriented] = field_grad(pts[oriented], pts[~oriented])
        while len(todo) > 0:
            scores = [(E[g] * pts[g, 3:]).sum(dim=-1).sum() for _, g in todo]
            j = torch.tensor(scores).abs().argmax().item()
            pid, rows = todo.pop(j)
            if trace is not None:
                trace.append(pid)
            if scores[j] < 0:
                pts[rows, 3:] *= -1
            oriented[rows] = True
            if diffuse:
                own = nothing.clone()
                own[rows] = True
                E[~own] = E[~own] + field_grad(pts[rows], pts[~own])
            else:
                E[~oriented] = E[~oriented] + field_grad(pts[rows], pts[~oriented])
        if diffuse:
            for _, rows in kept:
                pts[rows, 3:] = pts[rows, 3:] * _sign_of((E[rows] * pts[rows, 3:]).sum(dim=-1))[:, None]
        if weights is not None:
            pts[:, 3:] = pts[:, 3:] / weights[:, None]


def propagate_representatives(pts, reps, diffuse=False, weights=None, trace=None, stage_seconds=None):
    """field_utils.py:207-282.  reps = [(representative rows, remaining rows)] per patch: greedy loop over the
    representatives only, a flip takes the patch's remaining rows along, then ONE field of all representatives on
    all other points and a per-point sign.  stage_seconds (dict) receives the wall time of the two stages."""
    import time
    pts = pts.detach()
    with torch.no_grad():
        t0 = time.perf_counter()
        if weights is not None:
            weights = weights.clamp(0.1, 1)
            pts[:, 3:] = pts[:, 3:] * weights[:, None]
        todo = []
        E = torch.zeros_like(pts[:, :3])
        groups = []
        oriented = torch.zeros(len(pts)).bool()
        waiting = torch.zeros(len(pts)).bool()
        for k, (rep, rest) in enumerate(reps):
            todo.append((k, rep, rest))
            groups.append(rep)
            waiting[rep] = True
        pid, rows, _ = todo.pop(_flattest(pts, groups))
        if trace is not None:
            trace.append(pid)
        oriented[rows] = True
        waiting[rows] = False
        E[waiting] = field_grad(pts[oriented], pts[waiting])
        while len(todo) > 0:
            scores = [(E[rep] * pts[rep, 3:]).sum(dim=-1).sum() for _, rep, _ in todo]
            j = torch.tensor(scores).abs().argmax().item()
            pid, rows, rest = todo.pop(j)
            if trace is not None:
                trace.append(pid)
            if scores[j] < 0:
                pts[rows, 3:] *= -1
                pts[rest, 3:] *= -1
            oriented[rows] = True
            waiting[rows] = False
            if diffuse:
                others = torch.logical_or(oriented, waiting)
                others[rows] = False
                E[others] = E[others] + field_grad(pts[rows], pts[others])
            else:
                E[waiting] = E[waiting] + field_grad(pts[rows], pts[waiting])
        if diffuse:
            for rep, _ in reps:
                pts[rep, 3:] = pts[rep, 3:] * _sign_of((E[rep] * pts[rep, 3:]).sum(dim=-1))[:, None]
        t1 = time.perf_counter()
        E = field_grad(pts[oriented], pts[~oriented])
        dots = (E * pts[~oriented, 3:]).sum(dim=-1)
        pts[~oriented, 3:] = pts[~oriented, 3:] * _sign_of(dots)[:, None]
        if weights is not None:
            pts[:, 3:] = pts[:, 3:] / weights[:, None]
        if stage_seconds is not None:
            stage_seconds["propagation"] = t1 - t0
            stage_seconds["final_field"] = time.perf_counter() - t1
