"""Stand-ins for the three helpers of the reference's util.py that the dipole hot
path touches, used ONLY when the reference's own `util` module is not importable
(e.g. on a box without the reference checkout).  When `util` imports, the
drop-in field_utils uses it unchanged, as the reference does.

  gen_grid   util.py:26-36   (the lattice itself is generated in-kernel; this is for callers that want it)
  Transform  util.py:577-609 ('reg' variant: centre by the xyz mean, scale by the largest bounding-box extent)
"""
import torch


def gen_grid(n=10):
    i = torch.arange(0, n ** 3)
    coords = torch.stack([i // (n * n), (i // n) % n, i % n], dim=1).float()
    return (coords / n - 0.5) * 2


class Transform:
    def __init__(self, pc, ttype='reg'):
        if ttype != 'reg':
            raise NotImplementedError("only the 'reg' transform is provided outside the reference's util.py")
        xyz = pc[:, :3]
        self.center = xyz.mean(dim=0)
        self.scale = (xyz.max(dim=0)[0] - xyz.min(dim=0)[0]).max()

    def apply(self, pc):
        out = pc.clone()
        out[:, :3] = (out[:, :3] - self.center[None, :]) / self.scale
        return out

    def inverse(self, pc):
        out = pc.clone()
        out[:, :3] = out[:, :3] * self.scale + self.center[None, :]
        return out

    @staticmethod
    def trans(pc, ttype='reg'):
        t = Transform(pc, ttype=ttype)
        return t.apply(pc), t
