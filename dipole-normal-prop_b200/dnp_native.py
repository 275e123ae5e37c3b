"""ctypes binding of libdipole.so (include/dipole.h) for PyTorch tensors.

PyTorch is plumbing here: it owns device memory and streams.  All arithmetic of
the dipole hot path happens in the hand-written sm_100a kernels behind the C
ABI.  There is NO CPU fallback: importing this module without the built library
or calling it without a CUDA device raises.
"""
import ctypes
import os
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdipole.so")

DNP_OK = 0
EPI_STORE, EPI_ACCUM, EPI_SIGN_GE, EPI_SIGN_GT, EPI_NORMALIZE, EPI_DOT = range(6)

# every symbol include/dipole.h declares: (name, restype, argtypes)
_vp, _i64, _i32, _f32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_float
_sz = ctypes.c_size_t
SYMBOLS = [
    ("dnp_last_error", ctypes.c_char_p, []),
    ("dnp_version", _i32, []),
    ("dnp_build_info", ctypes.c_char_p, []),
    ("dnp_device_info", _i32, [ctypes.POINTER(_i32)] * 4),
    ("dnp_pairs_workspace_bytes", _sz, [_i64, _i64]),
    ("dnp_potential_f32", _i32, [_vp, _i64, _i64, _vp, _vp, _i64, _i64, _vp, _i32, _vp, _sz, _vp]),
    ("dnp_lattice_mean_potential_f32", _i32, [_vp, _i64, _i64, _i32, _vp, _vp, _vp, _sz, _vp]),
    ("dnp_field_grad_f32", _i32, [_vp, _i64, _i64, _vp, _i64, _i64, _f32, _vp, _i64, _i32, _vp, _sz, _vp]),
    ("dnp_field_apply_f32", _i32, [_vp, _i64, _i64, _vp, _vp, _i64, _i64, _vp, _vp, _f32, _i32,
                                   _vp, _i64, _vp, _i64, _vp, _vp, _sz, _vp]),
    ("dnp_edge_interaction_f32", _i32, [_vp, _i64, _i64, _vp, _i64, _i64, _f32, _vp, _vp, _sz, _vp]),
    ("dnp_edge_interactions_batched_f32", _i32, [_vp, _i64, _vp, _vp, _i64, _vp, _vp, _vp, _i64, _f32, _vp, _vp]),
    ("dnp_patch_interactions_f32", _i32, [_vp, _i64, _vp, _i64, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp]),
    ("dnp_field_sign_f32", _i32, [_vp, _i64, _vp, _i64, _vp, _i64, _i32, _vp]),
    ("dnp_patch_covariance_f32", _i32, [_vp, _i64, _vp, _vp, _i64, _vp, _vp]),
    ("dnp_points_max_n", _i64, []),
    ("dnp_points_workspace_bytes", _sz, [_i64]),
    ("dnp_points_propagate_f32", _i32, [_vp, _i64, _i64, _i64, _f32, _i32, _vp, _vp, _vp, _sz, _vp]),
    ("dnp_patch_propagate_workspace_bytes", _sz, [_i64, _i64, _i64, _i64]),
    ("dnp_patch_propagate_sharded_workspace_bytes", _sz, [_i64, _i64, _i64, _i64, _i64, _i64]),
    ("dnp_patch_propagate_f32", _i32, [_vp, _i64, _i64, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _i64, _i64, _f32, _i32,
                                       _vp, _i64, _vp, _vp, _vp, _sz, _vp]),
    ("dnp_points_batched_max_n", _i64, []),
    ("dnp_points_propagate_batched_f32", _i32, [_vp, _i64, _vp, _i64, _i64, _f32, _i32, _vp]),
    ("dnp_patch_lattice_sign_f32", _i32, [_vp, _i64, _vp, _i64, _i32, _vp, _i32, _vp]),
    ("dnp_xie_interaction_f32", _i32, [_vp, _i64, _i64, _vp, _i64, _i64, _f32, _vp, _vp, _i64, _vp]),
    ("dnp_xie_field_f32", _i32, [_vp, _i64, _i64, _vp, _i64, _i64, _f32, _vp, _vp, _i64, _vp]),
    ("dnp_xie_max_n", _i64, [_i32]),
    ("dnp_knn_radius2_f64", _i32, [_vp, _i64, _i64, _vp, _i64, _i64, _i32, _vp, _vp]),
    ("dnp_comm_unique_id", _i32, [_vp]),
    ("dnp_comm_init", _i32, [_vp, _i32, _i32]),
    ("dnp_comm_destroy", _i32, []),
    ("dnp_comm_rank", _i32, [_vp, _vp]),
    ("dnp_sharded_workspace_bytes", _sz, [_i64, _i64, _i64, _i64]),
    ("dnp_field_grad_sharded_f32", _i32, [_vp, _i64, _i64, _vp, _i64, _i64, _f32, _vp, _i64, _i32, _vp, _sz, _vp]),
    ("dnp_edge_interactions_sharded_f32", _i32, [_vp, _i64, _vp, _vp, _i64, _i64, _vp, _i64, _f32, _vp, _vp, _sz, _vp]),
    ("dnp_lattice_mean_potential_sharded_f32", _i32, [_vp, _i64, _i64, _i32, _vp, _vp, _sz, _vp]),
    ("dnp_lattice_mean_potential_noalloc_f32", _i32, [_vp, _i64, _i64, _i32, _vp, _vp]),
    ("dnp_patch_propagate_sharded_f32", _i32, [_vp, _i64, _i64, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _i64, _vp, _i64, _f32,
                                               _i32, _vp, _i64, _vp, _vp, _vp, _sz, _vp]),
    ("dnp_xie_workspace_bytes", _sz, [_i64, _i64]),
    ("dnp_xie_propagate_in_order_f32", _i32, [_vp, _i64, _i64, _vp, _i64, _f32, _vp, _i32, _vp, _vp, _vp, _sz, _vp]),
    ("dnp_voxel_partition_workspace_bytes", _sz, [_i64]),
    ("dnp_voxel_partition_f32", _i32, [_vp, _i64, _i64, _i32, ctypes.c_double, ctypes.c_double, _vp, _vp, _vp, _vp,
                                       _vp, _sz, _vp]),
    ("dnp_flip_solve_workspace_bytes", _sz, [_i64, _i32, _i32]),
    ("dnp_flip_solve_f64", _i32, [_vp, _vp, _i64, _i64, _i32, _i32, _i32, ctypes.c_uint64, _vp, _vp, _vp, _vp, _sz,
                                  _vp]),
    ("dnp_xyz_parse_workspace_bytes", _sz, [_i64, _i64]),
    ("dnp_xyz_parse_f32", _i32, [_vp, _i64, _i64, _vp, _i64, _vp, _vp, _sz, _vp]),
    ("dnp_xyz_format_workspace_bytes", _sz, [_i64]),
    ("dnp_xyz_format_f32", _i32, [_vp, _i64, _i32, _i64, _i64, _vp, _i64, _vp, _vp, _sz, _vp]),
    ("dnp_xyz_export_workspace_bytes", _sz, [_i64]),
    ("dnp_xyz_export_fd_f32", _i32, [_vp, _i64, _i32, _i64, _i64, _i32, _vp, _vp, _sz, _vp]),
    ("dnp_wire_decode_f64", _i32, [_vp, _i64, _vp, _i64, _vp]),
    ("dnp_wire_encode_f64", _i32, [_vp, _i64, _i64, _vp, _vp]),
    ("dnp_field_grad_host_f32", _i32, [_vp, _i64, _i64, _vp, _i64, _i64, _f32, _vp, _i32]),
    ("dnp_potential_host_f32", _i32, [_vp, _i64, _i64, _vp, _i64, _i64, _vp, _i32]),
    ("dnp_launch_count", _i64, []),
    ("dnp_profile_begin", _i32, []),
    ("dnp_profile_end", _i32, [ctypes.POINTER(_f32), _i32, ctypes.POINTER(_i32)]),
    ("dnp_microbench", _i32, [_i32, _i32, ctypes.POINTER(_f32), ctypes.POINTER(ctypes.c_double), _vp]),
]

_lib = None
_lock = threading.Lock()


class DipoleError(RuntimeError):
    pass


def source_hash():
    """The hash csrc/Makefile compiles into the library: sha256 over every .cu / .cuh / .h of csrc/ (sorted) and
    include/dipole.h; None when the sources are not next to the library (an installed copy)."""
    import glob
    import hashlib
    csrc = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc")
    header = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "dipole.h")
    files = sorted(glob.glob(os.path.join(csrc, "*.cu")) + glob.glob(os.path.join(csrc, "*.cuh")) +
                   glob.glob(os.path.join(csrc, "*.h")), key=os.path.basename)
    if not files or not os.path.exists(header):
        return None
    h = hashlib.sha256()
    for f in files + [header]:
        h.update(open(f, "rb").read())
    return h.hexdigest()[:16]


_BUILD_MARK = b"libdipole sm_100a sources="


def embedded_hash(path=LIB_PATH):
    """The source hash inside a built library (the literal dnp_build_info returns), read WITHOUT loading it."""
    import mmap
    with open(path, "rb") as f, mmap.mmap(f.fileno(), 0, access=mmap.ACCESS_READ) as m:
        i = m.find(_BUILD_MARK)
        if i < 0:
            return None
        return m[i + len(_BUILD_MARK):i + len(_BUILD_MARK) + 16].decode("ascii", "replace")


def _rebuild_if_stale():
    """A missing library, or one built from other sources than the tree next to it, is rebuilt in place before it is
    loaded when nvcc is on the machine (the GPU boxes carry the build container's toolchain); the processes of one
    job serialise on a lock file and re-check after acquiring it.  DNP_NO_AUTOBUILD=1 turns this off; without nvcc,
    or if the build fails, load() raises as before.  There is still no CPU fallback: this only ever runs `make`."""
    import fcntl
    import shutil
    import subprocess
    import sys
    want = source_hash()
    if want is None or os.environ.get("DNP_NO_AUTOBUILD") == "1":
        return
    csrc = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc")

    def stale():
        return not os.path.exists(LIB_PATH) or embedded_hash() != want
    if not stale() or not (shutil.which("nvcc") or os.path.exists("/usr/local/cuda/bin/nvcc")):
        return
    with open(os.path.join(csrc, ".build.lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        if stale():
            print(f"dnp_native: {LIB_PATH} does not match the sources next to it (tree {want}); running make",
                  file=sys.stderr, flush=True)
            env = dict(os.environ, PATH=os.environ.get("PATH", "") + ":/usr/local/cuda/bin")
            subprocess.run(["make", "-C", csrc, "-j8"], stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, env=env,
                           check=False)


def load():
    """Load libdipole.so (once).  Fails loudly if the CUDA extension is missing or was built from other sources."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                _rebuild_if_stale()
                if not os.path.exists(LIB_PATH):
                    raise ImportError(
                        f"{LIB_PATH} is missing: build it with `python __graft_entry__.py` "
                        "(or `make -C dipole-normal-prop_b200/csrc`). There is no CPU fallback.")
                lib = ctypes.CDLL(LIB_PATH)
                for name, restype, argtypes in SYMBOLS:
                    fn = getattr(lib, name)
                    fn.restype = restype
                    fn.argtypes = argtypes
                want = source_hash()
                info = lib.dnp_build_info().decode()
                if want is not None and f"sources={want}" not in info:
                    raise ImportError(f"{LIB_PATH} was built from other sources ({info}; this tree hashes to {want}): "
                                      "rebuild it with `make -C dipole-normal-prop_b200/csrc`")
                _lib = lib
    return _lib


def _check(rc):
    if rc != DNP_OK:
        raise DipoleError(f"libdipole error {rc}: {load().dnp_last_error().decode()}")


def require_cuda():
    if not torch.cuda.is_available():
        raise DipoleError("dipole-normal-prop_b200 needs a CUDA (sm_100a) device; there is no CPU fallback")


def _stream(t):
    return ctypes.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def _rows(t, min_cols, name):
    """Validate a [N, >=min_cols] float32 CUDA tensor; returns (tensor, ld) with unit column stride."""
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a torch.Tensor")
    if t.dtype != torch.float32:
        raise TypeError(f"{name} must be float32 (got {t.dtype}); the B200 kernels are FP32-only")
    if not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor")
    if t.dim() != 2 or t.shape[1] < min_cols:
        raise ValueError(f"{name} must be [N, >={min_cols}] (got {tuple(t.shape)})")
    if t.shape[0] > 0 and (t.stride(1) != 1 or t.stride(0) < t.shape[1]):
        t = t.contiguous()
    n = t.shape[0]
    ld = t.stride(0) if n > 1 else max(int(t.stride(0)) if n == 1 else 0, t.shape[1])
    return t, int(ld)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def _workspace(S, M, device):
    nbytes = load().dnp_pairs_workspace_bytes(int(S), int(M))
    return torch.empty(nbytes, dtype=torch.uint8, device=device), nbytes


def _idx(t, device, name):
    if t is None:
        return None
    if t.dtype != torch.int64 or t.device != device or not t.is_contiguous():
        t = t.to(device=device, dtype=torch.int64).contiguous()
    return t


def _mask(t, device, name):
    if t is None:
        return None
    if t.dtype == torch.bool:
        t = t.to(device).contiguous().view(torch.uint8)
    elif t.dtype != torch.uint8:
        raise TypeError(f"{name} must be bool or uint8")
    return t.to(device).contiguous()


# ------------------------------------------------------------------ leaf calls (device tensors)
def potential(sources, targets, out=None, accumulate=False, src_idx=None):
    lib = load()
    src, ld_s = _rows(sources, 6, "sources")
    tgt, ld_t = _rows(targets, 3, "targets")
    M = tgt.shape[0]
    if out is None:
        out = torch.empty(M, dtype=torch.float32, device=tgt.device)
    src_idx = _idx(src_idx, tgt.device, "src_idx")
    S = src.shape[0] if src_idx is None else src_idx.numel()
    if M == 0:
        return out
    with torch.cuda.device(tgt.device):
        ws, nbytes = _workspace(S, M, tgt.device)
        _check(lib.dnp_potential_f32(_ptr(src), S, ld_s, _ptr(src_idx), _ptr(tgt), M, ld_t, _ptr(out),
                                     int(accumulate), _ptr(ws), nbytes, _stream(tgt)))
    return out


def lattice_mean_potential(sources, n=10, want_phi=False):
    lib = load()
    src, ld_s = _rows(sources, 6, "pc")
    dev = src.device
    mean = torch.empty((), dtype=torch.float32, device=dev)
    phi = torch.empty(n ** 3, dtype=torch.float32, device=dev) if want_phi else None
    with torch.cuda.device(dev):
        ws, nbytes = _workspace(src.shape[0], n ** 3, dev)
        _check(lib.dnp_lattice_mean_potential_f32(_ptr(src), src.shape[0], ld_s, int(n), _ptr(phi), _ptr(mean),
                                                  _ptr(ws), nbytes, _stream(src)))
    return (mean, phi) if want_phi else mean


def field_apply(sources, targets, eps, mode, out=None, ld_out=None, normals=None, src_idx=None, tgt_idx=None,
                tgt_skip=None, nonfinite=None):
    """General all-pairs field call with a fused epilogue (dnp_field_apply_f32).

    `targets` is the FULL row array; tgt_idx / tgt_skip select rows of it, and
    `out` / `normals` are addressed by the same rows."""
    lib = load()
    src, ld_s = _rows(sources, 6, "sources")
    tgt, ld_t = _rows(targets, 3, "targets")
    dev = tgt.device
    src_idx = _idx(src_idx, dev, "src_idx")
    tgt_idx = _idx(tgt_idx, dev, "tgt_idx")
    tgt_skip = _mask(tgt_skip, dev, "tgt_skip")
    S = src.shape[0] if src_idx is None else src_idx.numel()
    M = tgt.shape[0] if tgt_idx is None else tgt_idx.numel()
    ld_n = 0
    if normals is not None:
        if normals.dtype != torch.float32 or not normals.is_cuda or normals.stride(-1) != 1:
            raise ValueError("normals must be a float32 CUDA view with unit column stride")
        ld_n = normals.stride(0) if normals.dim() == 2 else 3
    if out is not None and ld_out is None:
        ld_out = out.stride(0) if out.dim() == 2 else 1
    if M == 0:
        return out
    with torch.cuda.device(dev):
        ws, nbytes = _workspace(S, M, dev)
        _check(lib.dnp_field_apply_f32(_ptr(src), S, ld_s, _ptr(src_idx), _ptr(tgt), M, ld_t, _ptr(tgt_idx),
                                       _ptr(tgt_skip), float(eps), int(mode), _ptr(out), int(ld_out or 0),
                                       _ptr(normals), int(ld_n), _ptr(nonfinite), _ptr(ws), nbytes, _stream(tgt)))
    return out


def field_grad(sources, targets, eps=1e-5, out=None, accumulate=False, nonfinite=None):
    """nonfinite: optional int32 device counter, incremented by the number of Inf / NaN components zeroed."""
    tgt, _ = _rows(targets, 3, "targets")
    if out is None:
        out = torch.empty(tgt.shape[0], 3, dtype=torch.float32, device=tgt.device)
    return field_apply(sources, tgt, eps, EPI_ACCUM if accumulate else EPI_STORE, out=out, nonfinite=nonfinite)


def edge_interaction(S_pts, T_pts, eps=1e-5):
    """sum_t n_t . E_S(t) as a 0-d device tensor (dnp_edge_interaction_f32)."""
    lib = load()
    s, ld_s = _rows(S_pts, 6, "sources")
    t, ld_t = _rows(T_pts, 6, "means")
    out = torch.empty((), dtype=torch.float32, device=t.device)
    with torch.cuda.device(t.device):
        ws, nbytes = _workspace(s.shape[0], t.shape[0], t.device)
        _check(lib.dnp_edge_interaction_f32(_ptr(s), s.shape[0], ld_s, _ptr(t), t.shape[0], ld_t, float(eps),
                                            _ptr(out), _ptr(ws), nbytes, _stream(t)))
    return out


def edge_interactions_batched(pts, perm, offsets, edges, eps=1e-5, tgt_perm=None, tgt_offsets=None):
    """out[e] = sum_{t in target patch} n_t . E_{source patch}(t) for every edge, one launch.
    tgt_perm / tgt_offsets: optional second CSR for the target side (a rank's slice of every patch)."""
    lib = load()
    p, ld = _rows(pts, 6, "pts")
    dev = p.device
    perm = _idx(perm, dev, "perm")
    offsets = _idx(offsets, dev, "offsets")
    tgt_perm = _idx(tgt_perm, dev, "tgt_perm")
    tgt_offsets = _idx(tgt_offsets, dev, "tgt_offsets")
    edges = _idx(edges, dev, "edges")
    nE = edges.shape[0]
    out = torch.empty(nE, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _check(lib.dnp_edge_interactions_batched_f32(_ptr(p), ld, _ptr(perm), _ptr(offsets), offsets.numel() - 1,
                                                     _ptr(tgt_perm), _ptr(tgt_offsets), _ptr(edges), nE,
                                                     float(eps), _ptr(out), _stream(p)))
    return out


def patch_interactions(E, pts, perm, offsets, active=None, want_best=True):
    """out[p] = sum_{t in p} E_t . n_t for active patches.  With want_best also returns `best`, an int32[2]
    device tensor: best[0] = first abs-argmax over active patches, best[1] = float bits of its signed value
    (read both with ONE device->host copy: see best_to_host)."""
    lib = load()
    p, ld = _rows(pts, 6, "pts")
    e, ld_e = _rows(E, 3, "E")
    dev = p.device
    perm = _idx(perm, dev, "perm")
    offsets = _idx(offsets, dev, "offsets")
    P = offsets.numel() - 1
    out = torch.empty(P, dtype=torch.float32, device=dev)
    best = torch.empty(2, dtype=torch.int32, device=dev) if want_best else None
    active = _mask(active, dev, "active")
    normals = p[:, 3:]
    with torch.cuda.device(dev):
        _check(lib.dnp_patch_interactions_f32(
            _ptr(e), ld_e, _ptr(normals), ld, _ptr(perm), _ptr(offsets), P, _ptr(active), _ptr(out),
            _ptr(best), ctypes.c_void_p(best.data_ptr() + 4) if want_best else ctypes.c_void_p(0), _stream(p)))
    return out, best


def best_to_host(best):
    """(patch index, signed interaction) from the `best` tensor of patch_interactions: one D2H copy."""
    b = best.cpu()
    return int(b[0]), float(b[1:].view(torch.float32)[0])


def field_sign(E, pts, idx=None, strict=True):
    """In place: n_r *= sign(E_r . n_r) for r in idx (all rows if None); strict -> (> 0), else (>= 0)."""
    lib = load()
    if pts.dtype != torch.float32 or not pts.is_cuda or pts.dim() != 2 or pts.shape[1] < 6 or pts.stride(1) != 1:
        raise ValueError("pts must be a [N,6] float32 CUDA tensor with unit column stride")
    e, ld_e = _rows(E, 3, "E")
    dev = pts.device
    idx = _idx(idx, dev, "idx")
    n = pts.shape[0] if idx is None else idx.numel()
    if n == 0:
        return
    normals = pts[:, 3:]
    with torch.cuda.device(dev):
        _check(lib.dnp_field_sign_f32(_ptr(e), ld_e, _ptr(normals), int(pts.stride(0)), _ptr(idx), n,
                                      EPI_SIGN_GT if strict else EPI_SIGN_GE, _stream(pts)))


def patch_covariance(pts, perm, offsets):
    """[P,3,3] float64 covariance (about the mean, / patch size) of every CSR patch."""
    lib = load()
    p, ld = _rows(pts, 3, "pts")
    dev = p.device
    perm = _idx(perm, dev, "perm")
    offsets = _idx(offsets, dev, "offsets")
    P = offsets.numel() - 1
    cov = torch.empty(P, 3, 3, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _check(lib.dnp_patch_covariance_f32(_ptr(p), ld, _ptr(perm), _ptr(offsets), P, _ptr(cov), _stream(p)))
    return cov


def points_propagate(pts, diffuse=False, starting_point=0, eps=1e-6, want_trace=False):
    """In-place per-point greedy propagation on a [N,6] float32 CUDA tensor."""
    lib = load()
    if pts.dtype != torch.float32 or not pts.is_cuda or pts.dim() != 2 or pts.shape[1] != 6:
        raise ValueError("pts must be a [N,6] float32 CUDA tensor")
    if pts.stride(1) != 1:
        raise ValueError("pts must have unit column stride")
    N = pts.shape[0]
    order = torch.empty(N, dtype=torch.int32, device=pts.device) if want_trace else None
    flipped = torch.empty(N, dtype=torch.uint8, device=pts.device) if want_trace else None
    with torch.cuda.device(pts.device):
        # DNP_POINTS_FORCE_GRID (tests): run the cooperative-grid kernel at any N, so size its scratch
        nbytes = lib.dnp_points_workspace_bytes((1 << 20) if os.environ.get("DNP_POINTS_FORCE_GRID") else N)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=pts.device) if nbytes else None
        _check(lib.dnp_points_propagate_f32(_ptr(pts), N, pts.stride(0) if N > 1 else 6, int(starting_point),
                                            float(eps), int(bool(diffuse)), _ptr(order), _ptr(flipped),
                                            _ptr(ws), nbytes, _stream(pts)))
    return (order, flipped) if want_trace else None


# ------------------------------------------------------------------ "xie" interaction
def patch_propagate(pts, perm, offsets, start_patch, E, diffuse=False, eps=1e-5, rest_perm=None, rest_offsets=None,
                    tgt_idx=None, want_trace=False, offsets_host=None, owned=None):
    """The whole greedy patch loop of strongest_field_propagation{,_reps} in one native call
    (dnp_patch_propagate_f32), in place on pts [N, 6]; E [N, 3] must be zero and receives the final field.
    offsets / rest_offsets may be tensors on any device (a host copy is made: the launch shapes need the sizes);
    pass the device tensor as `offsets` AND its host copy as `offsets_host` to avoid any copy inside the call (needed
    to record the call into a CUDA graph).  Returns (order, values) as numpy arrays when want_trace.
    owned (uint8 [P], this rank's patches) selects the multi-GPU loop, dnp_patch_propagate_sharded_f32: tgt_idx must
    then be the rows of the owned patches (may be empty) and comm_init must have been called on every rank."""
    import numpy as np
    lib = load()
    p, ld = _rows(pts, 6, "pts")
    if p.data_ptr() != pts.data_ptr():
        raise ValueError("pts must have unit column stride (the call works in place)")
    e, ld_e = _rows(E, 3, "E")
    if e.data_ptr() != E.data_ptr():
        raise ValueError("E must have unit column stride")
    dev = p.device
    perm = _idx(perm, dev, "perm")
    off_dev = _idx(offsets, dev, "offsets")
    src_host = offsets if offsets_host is None else offsets_host
    off_host = np.ascontiguousarray(src_host.cpu().numpy() if torch.is_tensor(src_host) else src_host, dtype=np.int64)
    P = off_host.size - 1
    rest_perm = _idx(rest_perm, dev, "rest_perm")
    rest_host = None if rest_offsets is None else np.ascontiguousarray(rest_offsets.cpu().numpy(), dtype=np.int64)
    if (rest_perm is None) != (rest_host is None):
        raise ValueError("rest_perm and rest_offsets go together")
    tgt_idx = _idx(tgt_idx, dev, "tgt_idx")
    n_tgt = 0 if tgt_idx is None else tgt_idx.numel()
    order = np.zeros(max(P, 1), dtype=np.int32) if want_trace else None
    values = np.zeros(max(P, 1), dtype=np.float32) if want_trace else None
    max_patch = int((off_host[1:] - off_host[:-1]).max()) if P > 0 else 0
    N = p.shape[0]
    with torch.cuda.device(dev):
        def query():
            if owned is None:
                return lib.dnp_patch_propagate_workspace_bytes(N, P, max_patch, n_tgt)
            # the slab schedule takes ALL patch rows as targets on every rank and holds one slab per OWNED patch
            return lib.dnp_patch_propagate_sharded_workspace_bytes(N, P, max_patch, n_tgt, int(perm.numel()),
                                                                   int(owned.sum().item()))
        nbytes = query()
        slab_was = os.environ.get("DNP_SLAB")
        try:
            ws = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=dev)
        except torch.OutOfMemoryError:
            # the slab schedule's per-patch field slabs (P x T x 12 bytes) do not fit next to what the caller holds:
            # same decisions through the per-step schedules and their small workspace
            # (not for the sharded loop: every rank must take the same schedule, or their collectives do not match)
            if slab_was == "0" or owned is not None:
                raise
            os.environ["DNP_SLAB"] = "0"
            nbytes = query()
            ws = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=dev)
        head = (_ptr(p), N, ld, _ptr(perm), _ptr(off_dev), ctypes.c_void_p(off_host.ctypes.data), P, _ptr(rest_perm),
                ctypes.c_void_p(rest_host.ctypes.data if rest_host is not None else 0), _ptr(tgt_idx), n_tgt)
        tail = (int(start_patch), float(eps), int(bool(diffuse)), _ptr(e), ld_e,
                ctypes.c_void_p(order.ctypes.data if want_trace else 0),
                ctypes.c_void_p(values.ctypes.data if want_trace else 0), _ptr(ws), nbytes, _stream(p))
        try:
            if owned is None:
                _check(lib.dnp_patch_propagate_f32(*head, *tail))
            else:
                if tgt_idx is None:
                    raise ValueError("the sharded loop needs tgt_idx (the rows of this rank's patches)")
                owned = owned.to(device=dev, dtype=torch.uint8).contiguous()
                if owned.numel() != P:
                    raise ValueError("owned must have one entry per patch")
                _check(lib.dnp_patch_propagate_sharded_f32(*head, _ptr(owned), *tail))
        finally:
            if os.environ.get("DNP_SLAB") != slab_was:
                if slab_was is None:
                    del os.environ["DNP_SLAB"]
                else:
                    os.environ["DNP_SLAB"] = slab_was
    return (order[:P], values[:P]) if want_trace else None


def points_propagate_batched(pts, offsets, max_patch, diffuse=False, eps=1e-6):
    """strongest_field_propagation_points on every patch of a CSR in ONE launch (a cluster per patch), in place.
    pts: [N, 6] float32 CUDA rows stored patch by patch; offsets: int64 [P + 1]."""
    lib = load()
    p, ld = _rows(pts, 6, "pts")
    if p.data_ptr() != pts.data_ptr():
        raise ValueError("pts must have unit column stride (the call works in place)")
    off = _idx(offsets, p.device, "offsets")
    with torch.cuda.device(p.device):
        _check(lib.dnp_points_propagate_batched_f32(_ptr(p), ld, _ptr(off), off.numel() - 1, int(max_patch),
                                                    float(eps), int(bool(diffuse)), _stream(p)))
    return pts


def patch_lattice_sign(pts, offsets, n=10, flip=True):
    """Mean lattice potential of every patch (float32 [P]); with flip, patches with a negative mean are negated."""
    lib = load()
    p, ld = _rows(pts, 6, "pts")
    if p.data_ptr() != pts.data_ptr():
        raise ValueError("pts must have unit column stride (the call works in place)")
    off = _idx(offsets, p.device, "offsets")
    P = off.numel() - 1
    mean = torch.zeros(P, dtype=torch.float32, device=p.device)
    with torch.cuda.device(p.device):
        _check(lib.dnp_patch_lattice_sign_f32(_ptr(p), ld, _ptr(off), P, int(n), _ptr(mean), int(bool(flip)),
                                              _stream(p)))
    return mean


def knn_radius2(sources, targets, k):
    """Squared distance (float64) from every source to its k-th nearest target: the kNN mask of
    field_utils.py:447-455 as one radius per source (dnp_knn_radius2_f64; exact differences of the float coordinates in
    double, the same expression the xie kernels test)."""
    lib = load()
    src, ld_s = _rows(sources, 3, "sources")
    tgt, ld_t = _rows(targets, 3, "targets")
    k = min(int(k), tgt.shape[0])
    out = torch.empty(src.shape[0], dtype=torch.float64, device=src.device)
    if src.shape[0] == 0:
        return out
    with torch.cuda.device(src.device):
        _check(lib.dnp_knn_radius2_f64(_ptr(src), src.shape[0], ld_s, _ptr(tgt), tgt.shape[0], ld_t, k, _ptr(out),
                                       _stream(src)))
    return out


def xie_matrix(sources, targets, C=3.0, radius2=None, field=False):
    """[M, S] interaction matrix (field=False, targets need normals) or [M, S, 3] field (field=True)."""
    lib = load()
    src, ld_s = _rows(sources, 6, "source")
    tgt, ld_t = _rows(targets, 3 if field else 6, "target")
    S, M = src.shape[0], tgt.shape[0]
    out = torch.empty((M, S, 3) if field else (M, S), dtype=torch.float32, device=tgt.device)
    if S == 0 or M == 0:
        return out
    if radius2 is not None:
        radius2 = radius2.to(device=tgt.device, dtype=torch.float64).contiguous()
    fn = lib.dnp_xie_field_f32 if field else lib.dnp_xie_interaction_f32
    with torch.cuda.device(tgt.device):
        _check(fn(_ptr(src), S, ld_s, _ptr(tgt), M, ld_t, float(C), _ptr(radius2), _ptr(out),
                  3 * S if field else S, _stream(tgt)))
    return out


def xie_propagate_in_order(pts, order, C=3.0, radius2=None, diffuse=False):
    """(flips[T, N] bool, interactions[T, N] float32) of the in-order propagation, one launch."""
    lib = load()
    p, ld = _rows(pts, 6, "pts")
    dev = p.device
    order = _idx(torch.as_tensor(order), dev, "order")
    if order.dim() != 2 or order.shape[1] != p.shape[0]:
        raise ValueError(f"order must be [T, N={p.shape[0]}] (got {tuple(order.shape)})")
    T, N = order.shape
    if N > 0 and (int(order.min()) < 0 or int(order.max()) >= N):
        raise ValueError("order holds indices outside [0, N)")
    flips = torch.zeros(T, N, dtype=torch.uint8, device=dev)
    inter = torch.zeros(T, N, dtype=torch.float32, device=dev)
    if radius2 is not None:
        radius2 = radius2.to(device=dev, dtype=torch.float64).contiguous()
    with torch.cuda.device(dev):
        nbytes = lib.dnp_xie_workspace_bytes(N, T) if diffuse else 0
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev) if nbytes else None
        _check(lib.dnp_xie_propagate_in_order_f32(_ptr(p), N, ld, _ptr(order), T, float(C), _ptr(radius2),
                                                  int(bool(diffuse)), _ptr(flips), _ptr(inter), _ptr(ws), nbytes,
                                                  _stream(p)))
    return flips.bool(), inter


# ------------------------------------------------------------------ voxel partition
def voxel_partition(pts, n_part, ranges=(-1.5, 1.5)):
    """CSR of the non-empty voxel cells of util._divide_pc: (perm[n_in] int64, offsets[n_cells+1] int64,
    ijk[n_cells, 3] int32).  One device->host read of two counters sizes the outputs."""
    lib = load()
    p, ld = _rows(pts, 3, "pc_in")
    dev = p.device
    N = p.shape[0]
    max_cells = min(N, (int(n_part) + 1) ** 3)
    perm = torch.empty(N, dtype=torch.int64, device=dev)
    offsets = torch.zeros(max_cells + 1, dtype=torch.int64, device=dev)
    ijk = torch.empty(max_cells, 3, dtype=torch.int32, device=dev)
    counts = torch.zeros(2, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        nbytes = lib.dnp_voxel_partition_workspace_bytes(N)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        _check(lib.dnp_voxel_partition_f32(_ptr(p), N, ld, int(n_part), float(ranges[0]), float(ranges[1]),
                                           _ptr(perm), _ptr(offsets), _ptr(ijk), _ptr(counts), _ptr(ws), nbytes,
                                           _stream(p)))
    n_cells, n_in = counts.tolist()
    return perm[:n_in], offsets[:n_cells + 1], ijk[:n_cells]


# ------------------------------------------------------------------ patch-graph flip solve
FLIP_EXACT_MAX = 36      # exhaustive up to this many patches (2^35 states, a few ms); the ABI allows 40
FLIP_BEST_FOUND, FLIP_EXHAUSTIVE, FLIP_BOUND_REACHED = 0, 1, 2


def flip_solve(A, B, exact_max_n=FLIP_EXACT_MAX, replicas=None, sweeps=None, seed=0):
    """argmax over x in {0,1}^n of graph.cal_loss(x, A, B) (graph.py:431-470).  A, B: [n, n] float64 CUDA
    tensors.  Returns device tensors (x uint8 [n], obj float64 [2] = (objective, bound), status int32 [1]).
    Annealing budget (n > exact_max_n) defaults to 2368 chains x 512 sweeps up to 256 patches and shrinks with
    n^2 beyond (a sweep costs O(n^2 / 32) per chain), so that a solve stays in the 10-100 ms range."""
    n_ = int(A.shape[0]) if hasattr(A, "shape") and len(A.shape) else 0
    if sweeps is None:
        sweeps = max(32, min(512, int(512 * (256.0 / max(n_, 1)) ** 2)))
    if replicas is None:
        replicas = 2368 if n_ <= 1024 else 592
    lib = load()
    for name, m in (("A", A), ("B", B)):
        if not isinstance(m, torch.Tensor) or m.dtype != torch.float64 or not m.is_cuda:
            raise TypeError(f"{name} must be a float64 CUDA tensor")
        if m.dim() != 2 or m.shape[0] != m.shape[1] or m.shape != A.shape:
            raise ValueError("A and B must be square matrices of the same shape")
    n = A.shape[0]
    dev = A.device
    x = torch.zeros(n, dtype=torch.uint8, device=dev)
    obj = torch.zeros(2, dtype=torch.float64, device=dev)
    status = torch.full((1,), FLIP_EXHAUSTIVE, dtype=torch.int32, device=dev)
    if n == 0:
        return x, obj, status
    A = A.contiguous()
    B = B.contiguous()
    with torch.cuda.device(dev):
        nbytes = lib.dnp_flip_solve_workspace_bytes(n, int(exact_max_n), int(replicas))
        if nbytes == 0:
            raise DipoleError(f"flip solve supports 1..4096 patches (got {n})")
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        _check(lib.dnp_flip_solve_f64(_ptr(A), _ptr(B), n, n, int(exact_max_n), int(replicas), int(sweeps),
                                      int(seed) & (2 ** 64 - 1), _ptr(x), _ptr(obj), _ptr(status), _ptr(ws), nbytes,
                                      _stream(A)))
    return x, obj, status


# ------------------------------------------------------------------ text IO + wire format
def xyz_parse(text_bytes, device=None):
    """util.xyz2tensor's parse on the device.  text_bytes: bytes / bytearray / memoryview / uint8 tensor with the
    file's content.  Returns (rows [n, 6] float32 CUDA tensor, column mask, first unparsable line or -1, number of
    lines).  Host side: ONE copy of the text into a reused pinned buffer, one asynchronous upload; the lines are
    counted on the device (no Python pass over the text)."""
    import warnings
    lib = load()
    require_cuda()
    dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
    if isinstance(text_bytes, torch.Tensor):
        txt = text_bytes.to(dev, non_blocking=True)
    elif len(text_bytes) == 0:
        txt = torch.empty(0, dtype=torch.uint8, device=dev)
    else:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")                 # read-only buffer: it is only read
            view = torch.frombuffer(memoryview(text_bytes), dtype=torch.uint8)
        txt = view.to(dev)
    n_bytes = txt.numel()
    with torch.cuda.device(dev):
        n_lines = int((txt == 10).sum().item()) + 1
        out = torch.empty(n_lines, 6, dtype=torch.float32, device=dev)
        counts = torch.empty(4, dtype=torch.int64, device=dev)
        nbytes = lib.dnp_xyz_parse_workspace_bytes(n_bytes, n_lines)
        if nbytes == 0:
            raise DipoleError("xyz parse: at most 2 GiB of text per call")
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        _check(lib.dnp_xyz_parse_f32(_ptr(txt), n_bytes, n_lines, _ptr(out), n_lines, _ptr(counts), _ptr(ws), nbytes,
                                     _stream(txt)))
    rows, mask, bad, lines = counts.tolist()
    return out[:rows], mask, bad, lines


def xyz_format(pc_rows, as_view=False):
    """util.export_pc's text for a [N, C] float32 CUDA tensor (any strides): returns the bytes, or with as_view a
    memoryview of the downloaded buffer (what dnp_util.export_pc writes to the file without another copy)."""
    lib = load()
    if pc_rows.dtype != torch.float32 or not pc_rows.is_cuda or pc_rows.dim() != 2:
        raise TypeError("xyz_format needs a 2-D float32 CUDA tensor")
    n, c = pc_rows.shape
    if n == 0 or c == 0:
        return b""
    dev = pc_rows.device
    cap = 25 * n * c
    out = torch.empty(cap, dtype=torch.uint8, device=dev)
    n_out = torch.zeros(1, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        nbytes = lib.dnp_xyz_format_workspace_bytes(n * c)
        if nbytes == 0:
            raise DipoleError("xyz format: too many values for one call")
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        _check(lib.dnp_xyz_format_f32(_ptr(pc_rows), n, c, pc_rows.stride(0), pc_rows.stride(1), _ptr(out), cap,
                                      _ptr(n_out), _ptr(ws), nbytes, _stream(pc_rows)))
        host = out[:int(n_out.item())].cpu().numpy()
    view = memoryview(host)
    return view if as_view else view.tobytes()


def xyz_export_fd(pc_rows, fd):
    """util.export_pc's text for a [N, C] float32 CUDA tensor (any strides), written to the open file descriptor `fd`
    by the library (device -> two pinned staging buffers -> write(2), no host copy of the whole text).  Returns the
    number of bytes written."""
    lib = load()
    if pc_rows.dtype != torch.float32 or not pc_rows.is_cuda or pc_rows.dim() != 2:
        raise TypeError("xyz_export_fd needs a 2-D float32 CUDA tensor")
    n, c = pc_rows.shape
    if n == 0 or c == 0:
        return 0
    dev = pc_rows.device
    written = ctypes.c_int64(0)
    with torch.cuda.device(dev):
        nbytes = lib.dnp_xyz_export_workspace_bytes(n * c)
        if nbytes == 0:
            raise DipoleError("xyz export: too many values for one call")
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        _check(lib.dnp_xyz_export_fd_f32(_ptr(pc_rows), n, c, pc_rows.stride(0), pc_rows.stride(1), int(fd),
                                         ctypes.byref(written), _ptr(ws), nbytes, _stream(pc_rows)))
    return int(written.value)


def wire_decode(xyz_f64, device=None):
    """N x 3 float64 request payload (numpy array, bytes or tensor) -> [N, 6] float32 CUDA rows, zero normals."""
    import numpy as np
    lib = load()
    require_cuda()
    if isinstance(xyz_f64, (bytes, bytearray, memoryview)):
        xyz_f64 = np.frombuffer(xyz_f64, dtype=np.float64).copy()
    t = torch.as_tensor(xyz_f64, dtype=torch.float64).reshape(-1, 3)
    dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
    t = t.to(dev).contiguous()
    out = torch.empty(t.shape[0], 6, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _check(lib.dnp_wire_decode_f64(_ptr(t), t.shape[0], _ptr(out), 6, _stream(t)))
    return out


def wire_encode(pc):
    """[N, 6] float32 CUDA rows -> the reply payload: N x 6 float64 as bytes."""
    lib = load()
    p, ld = _rows(pc, 6, "pc")
    out = torch.empty(p.shape[0], 6, dtype=torch.float64, device=p.device)
    with torch.cuda.device(p.device):
        _check(lib.dnp_wire_encode_f64(_ptr(p), p.shape[0], ld, _ptr(out), _stream(p)))
    return out.cpu().numpy().tobytes()


# ------------------------------------------------------------------ host-buffer calls (CPU tensors in, CPU tensors out)
def _host_rows(t, min_cols, name):
    if t.dtype != torch.float32:
        raise TypeError(f"{name} must be float32 (got {t.dtype})")
    if t.dim() != 2 or t.shape[1] < min_cols:
        raise ValueError(f"{name} must be [N, >={min_cols}]")
    if t.shape[0] > 0 and (t.stride(1) != 1 or t.stride(0) < t.shape[1]):
        t = t.contiguous()
    return t, int(t.stride(0) if t.shape[0] > 1 else t.shape[1])


def field_grad_host(sources, targets, eps=1e-5, device=None, out=None):
    require_cuda()
    lib = load()
    s, ld_s = _host_rows(sources, 6, "sources")
    t, ld_t = _host_rows(targets, 3, "means")
    if out is None:
        out = torch.empty(t.shape[0], 3, dtype=torch.float32)
    dev = torch.cuda.current_device() if device is None else int(device)
    _check(lib.dnp_field_grad_host_f32(_ptr(s), s.shape[0], ld_s, _ptr(t), t.shape[0], ld_t, float(eps), _ptr(out), dev))
    return out


def potential_host(sources, targets, device=None, out=None):
    require_cuda()
    lib = load()
    s, ld_s = _host_rows(sources, 6, "sources")
    t, ld_t = _host_rows(targets, 3, "means")
    if out is None:
        out = torch.empty(t.shape[0], dtype=torch.float32)
    dev = torch.cuda.current_device() if device is None else int(device)
    _check(lib.dnp_potential_host_f32(_ptr(s), s.shape[0], ld_s, _ptr(t), t.shape[0], ld_t, _ptr(out), dev))
    return out


# ------------------------------------------------------------------ multi-GPU entry points of the C ABI (dnp_comm.cu)
def comm_unique_id():
    """128-byte NCCL id (uint8 CPU tensor) created by the calling rank; hand it to the other ranks by any means."""
    ident = torch.zeros(128, dtype=torch.uint8)
    _check(load().dnp_comm_unique_id(_ptr(ident)))
    return ident


def comm_init(ident, rank, world):
    ident = ident.cpu().contiguous()
    _check(load().dnp_comm_init(_ptr(ident), int(rank), int(world)))


def comm_destroy():
    _check(load().dnp_comm_destroy())


def field_grad_sharded(sources, targets, eps=1e-5, gather=True):
    """dnp_field_grad_sharded_f32: this rank's block of E (all blocks when gather) -- shard kernels + NCCL exchange on
    the caller's stream, no torch.distributed involved."""
    lib = load()
    src, ld_s = _rows(sources, 6, "sources")
    tgt, ld_t = _rows(targets, 3, "targets")
    S, M = src.shape[0], tgt.shape[0]
    E = torch.zeros(M, 3, dtype=torch.float32, device=tgt.device)
    with torch.cuda.device(tgt.device):
        nbytes = lib.dnp_sharded_workspace_bytes(S, M, 0, 0)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=tgt.device)
        _check(lib.dnp_field_grad_sharded_f32(_ptr(src), S, ld_s, _ptr(tgt), M, ld_t, float(eps), _ptr(E), 3,
                                              int(bool(gather)), _ptr(ws), nbytes, _stream(tgt)))
    return E


def edge_interactions_sharded(pts, perm, offsets, edges, eps=1e-5):
    """dnp_edge_interactions_sharded_f32: per-edge sums with the target patches sliced over the ranks + ONE all-reduce."""
    lib = load()
    p, ld = _rows(pts, 6, "pts")
    dev = p.device
    perm, offsets, edges = _idx(perm, dev, "perm"), _idx(offsets, dev, "offsets"), _idx(edges, dev, "edges")
    P, rows = offsets.numel() - 1, perm.numel()
    out = torch.empty(edges.shape[0], dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        nbytes = lib.dnp_sharded_workspace_bytes(1, 1, P, rows)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        _check(lib.dnp_edge_interactions_sharded_f32(_ptr(p), ld, _ptr(perm), _ptr(offsets), P, rows, _ptr(edges),
                                                     edges.shape[0], float(eps), _ptr(out), _ptr(ws), nbytes, _stream(p)))
    return out


def lattice_mean_potential_sharded(sources, n=10):
    lib = load()
    src, ld_s = _rows(sources, 6, "pc")
    mean = torch.empty((), dtype=torch.float32, device=src.device)
    with torch.cuda.device(src.device):
        nbytes = lib.dnp_sharded_workspace_bytes(src.shape[0], n ** 3, 0, 0)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=src.device)
        _check(lib.dnp_lattice_mean_potential_sharded_f32(_ptr(src), src.shape[0], ld_s, int(n), _ptr(mean), _ptr(ws),
                                                          nbytes, _stream(src)))
    return mean


def lattice_mean_potential_noalloc(sources, n=10, out=None):
    """dnp_lattice_mean_potential_noalloc_f32: the lattice test's mean potential with the library's static device
    scratch instead of a workspace -- the CALL allocates nothing (pass `out`, a float32 CUDA scalar, to keep this
    wrapper allocation-free too).  Calls on one device share the scratch: do not overlap them on the GPU."""
    lib = load()
    src, ld_s = _rows(sources, 6, "pc")
    mean = out if out is not None else torch.empty((), dtype=torch.float32, device=src.device)
    if mean.dtype != torch.float32 or mean.device != src.device or mean.numel() != 1:
        raise TypeError("out must be a float32 scalar on the sources' device")
    with torch.cuda.device(src.device):
        _check(lib.dnp_lattice_mean_potential_noalloc_f32(_ptr(src), src.shape[0], ld_s, int(n), _ptr(mean),
                                                          _stream(src)))
    return mean


# ------------------------------------------------------------------ measurement helpers
def launch_count():
    return int(load().dnp_launch_count())


def profile_begin():
    _check(load().dnp_profile_begin())


def profile_end(capacity=4096):
    """Durations (ms) of the all-pairs kernels this thread launched since profile_begin()."""
    buf = (ctypes.c_float * capacity)()
    n = ctypes.c_int(0)
    _check(load().dnp_profile_end(buf, capacity, ctypes.byref(n)))
    return [buf[i] for i in range(min(n.value, capacity))]


def device_info():
    vals = [ctypes.c_int(0) for _ in range(4)]
    _check(load().dnp_device_info(*[ctypes.byref(v) for v in vals]))
    return {"sm_count": vals[0].value, "sm_clock_khz": vals[1].value, "cc": (vals[2].value, vals[3].value)}


def microbench(kind, iters=2000):
    ms, ops = ctypes.c_float(0), ctypes.c_double(0)
    _check(load().dnp_microbench(int(kind), int(iters), ctypes.byref(ms), ctypes.byref(ops),
                                 ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return ms.value, ops.value
