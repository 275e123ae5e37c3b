"""Torch (CPU) port of the reference's leaf evaluation, op for op.  TEST INFRASTRUCTURE ONLY
(see oracle/__init__.py): it is the CPU baseline that bench.py times beside the GPU numbers
(`cpu_baseline`, `--impl reference`) and a second checker in tests/.

Why a port: the reference is pure Python over ATen ops and cannot travel to the GPU box
(/root/reference does not exist there).  What the reference spends its time on is not the ~35 flops
per pair but the ~20 full-size [S, M, 3] temporaries, the four boolean-mask gathers/scatters and the
recursive chunking around them, so the port keeps exactly that op chain
and chunking: field_utils.py:23-44,73-94 (recursion), :46-55 (potential leaf), :96-116 (field leaf).
Parity status: PINNED -- tests/test_oracle.py checks it against tests/golden/leaf.npz (bit-for-bit for
the float32 outputs, since the op sequence is the reference's).
"""
import torch


def _halves(n):
    h = int(n / 2)
    return slice(0, h), slice(h, n)


def _chunked(leaf, sources, means, eps, max_pts):
    """field_utils.py:23-44 / :73-94: split the larger side in half until both sides fit max_pts;
    source halves are added, target halves concatenated."""
    S, M = sources.shape[0], means.shape[0]
    split_sources = S > max_pts and (M <= max_pts or S > M)
    split_means = (not split_sources) and M > max_pts
    if split_sources:
        a, b = _halves(S)
        return _chunked(leaf, sources[a], means, eps, max_pts) + _chunked(leaf, sources[b], means, eps, max_pts)
    if split_means:
        a, b = _halves(M)
        return torch.cat([_chunked(leaf, sources, means[a], eps, max_pts),
                          _chunked(leaf, sources, means[b], eps, max_pts)], dim=0)
    return leaf(sources, means, eps)


def _potential_leaf(sources, means, eps):
    moments = sources[:, 3:]
    disp = sources[:, None, :3] - means[None, :, :3]                     # [S, M, 3]
    phi = (moments[:, None, :] * disp).sum(dim=-1)
    phi = phi / (disp.norm(dim=-1) ** 3)
    total = phi.sum(dim=0)
    total[total.isinf()] = 0
    total[total.isnan()] = 0
    return total


def _field_leaf(sources, means, eps):
    moments = sources[:, 3:]
    disp = sources[:, None, :3] - means[None, :, :3]                     # [S, M, 3]
    coincident = disp.norm(dim=-1) == 0
    unit = disp.clone()
    unit[~coincident] = disp[~coincident] / disp[~coincident].norm(dim=-1)[:, None]
    unit[coincident] = 0
    term = 3 * (moments[:, None, :] * unit).sum(dim=-1)[:, :, None] * unit - moments[:, None, :]
    term[coincident] = 0
    term = term / (disp.norm(dim=-1) ** 3 + eps)[:, :, None]
    total = term.sum(dim=0) * -1
    if total.isinf().any() or total.isnan().any():
        pass                                                             # the reference prints a count here
    total[total.isinf()] = 0
    total[total.isnan()] = 0
    return total


def potential(sources, means, eps=1e-5, recursive=True, max_pts=15000):
    if recursive:
        return _chunked(_potential_leaf, sources, means, eps, max_pts)
    return _potential_leaf(sources, means, eps)


def field_grad(sources, means, eps=1e-5, recursive=True, max_pts=15000):
    if recursive:
        return _chunked(_field_leaf, sources, means, eps, max_pts)
    return _field_leaf(sources, means, eps)
