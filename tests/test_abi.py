"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports exactly what
include/dipole.h declares, the ctypes table covers it, the Python mirror keeps the reference's
signatures, and nothing falls back to the CPU when there is no GPU.  No compute calls here."""
import ctypes
import inspect
import json
import os
import re
import subprocess

import pytest
import torch

from conftest import GOLDEN, PKG, ROOT

HEADER = os.path.join(ROOT, "include", "dipole.h")
LIB = os.path.join(PKG, "libdipole.so")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"DNP_API\s+[\w\s\*]+?\b(dnp_\w+)\s*\(", text)))


def exported_symbols():
    out = subprocess.check_output(["nm", "-D", "--defined-only", LIB], text=True)
    return sorted(line.split()[-1] for line in out.splitlines() if " T " in line and "dnp_" in line)


def test_library_is_built():
    assert os.path.exists(LIB), "build libdipole.so first: python -c 'import __graft_entry__ as g; g.build()'"


def test_library_was_built_from_this_source_tree(monkeypatch):
    """Build provenance: csrc/Makefile compiles a hash of every source into the library; dnp_native.load() refuses a
    prebuilt libdipole.so whose hash differs from the tree it sits in (the .so is git-ignored but travels to the GPU
    box with the snapshot)."""
    import dnp_native
    info = dnp_native.load().dnp_build_info().decode()
    assert "sm_100a" in info and f"sources={dnp_native.source_hash()}" in info, info
    assert dnp_native.embedded_hash() == dnp_native.source_hash()      # readable without loading the library
    # a stale library: load() first tries `make` in csrc/ (nvcc is on the build container and on the GPU boxes) ...
    made = []
    monkeypatch.setattr(subprocess, "run", lambda cmd, **kw: made.append(cmd))
    monkeypatch.setattr(dnp_native, "_lib", None)
    monkeypatch.setattr(dnp_native, "source_hash", lambda: "0123456789abcdef")
    with pytest.raises(ImportError, match="built from other sources"):      # ... and refuses what is still stale
        dnp_native.load()
    assert made and made[0][0] == "make"
    made.clear()
    monkeypatch.setenv("DNP_NO_AUTOBUILD", "1")
    with pytest.raises(ImportError, match="built from other sources"):
        dnp_native.load()
    assert not made


def test_header_and_exports_agree():
    decl, exp = declared_symbols(), exported_symbols()
    assert len(decl) >= 15
    assert decl == exp, f"header-only: {set(decl) - set(exp)}; undeclared exports: {set(exp) - set(decl)}"


def test_ctypes_table_covers_header():
    import dnp_native
    assert sorted(name for name, _, _ in dnp_native.SYMBOLS) == declared_symbols()
    lib = dnp_native.load()                       # dlopen + every symbol resolves
    assert lib.dnp_version() == 100
    assert lib.dnp_last_error() is not None
    assert lib.dnp_points_max_n() > 90000


def test_only_c_types_cross_the_boundary():
    text = open(HEADER).read()
    assert 'extern "C"' in text
    code = re.sub(r"/\*.*?\*/", "", text, flags=re.S)       # declarations only, comments stripped
    for banned in ("torch", "at::", "std::", "Tensor", "#include <cuda", "template"):
        assert banned not in code, banned


def test_every_entry_point_cites_the_reference():
    """Each block of the header names the reference lines it replaces (file:line)."""
    text = open(HEADER).read()
    assert len(re.findall(r"(field_utils|util|graph_dipole)\.py:\d+", text)) >= 15


def test_python_mirror_keeps_reference_signatures():
    import field_utils
    want = json.load(open(os.path.join(GOLDEN, "signatures.json")))
    for name, params in want["field_utils"].items():
        fn = getattr(field_utils, name)
        got = [[p.name, None if p.default is inspect.Parameter.empty else repr(p.default)]
               for p in inspect.signature(fn).parameters.values()]
        assert got == params, name
    for name in want["reexports"]:                # `from field_utils import *` users rely on these
        assert hasattr(field_utils, name)


def test_thin_callers_keep_reference_signatures():
    import graph_dipole
    assert list(inspect.signature(graph_dipole.propagate_points_file).parameters) == \
        ["input_pc", "propagate_func", "args", "kwargs"]
    sig = inspect.signature(graph_dipole.st_propagation_points_file)
    assert list(sig.parameters) == ["input_pc", "verbose"] and sig.parameters["verbose"].default is True


def test_product_never_imports_the_oracle():
    """No import, dlopen or link of anything under oracle/ from the shipped package."""
    pat = re.compile(r"^\s*(import|from)\s+\S*oracle|liboracle|dipole_oracle|orc_\w+\(", re.M)
    for fn in os.listdir(PKG):
        if fn.endswith(".py"):
            assert not pat.search(open(os.path.join(PKG, fn)).read()), fn
    for fn in os.listdir(os.path.join(PKG, "csrc")):
        if fn.endswith((".cu", ".cuh", ".h", "Makefile")):
            assert not pat.search(open(os.path.join(PKG, "csrc", fn)).read()), fn
    needed = subprocess.check_output(["readelf", "-d", LIB], text=True)
    assert "oracle" not in needed


def test_dtype_and_shape_errors_are_the_wrappers():
    import field_utils
    a = torch.zeros(4, 6, dtype=torch.float64)
    with pytest.raises(TypeError):
        field_utils.field_grad(a, a)
    with pytest.raises(TypeError):
        field_utils.potential(a.numpy(), a.numpy())


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks behaviour on a box WITHOUT a GPU")
def test_no_cpu_fallback():
    """Without a CUDA device the product path fails loudly instead of computing on the host."""
    import dnp_native
    import field_utils
    a = torch.rand(8, 6)
    for call in (lambda: field_utils.field_grad(a, a), lambda: field_utils.potential(a, a),
                 lambda: field_utils.measure_mean_potential(a),
                 lambda: field_utils.strongest_field_propagation_points(a.clone()),
                 lambda: field_utils.field_edge_calculator(a, a),
                 lambda: field_utils.reference_field(a, a.clone())):
        with pytest.raises(dnp_native.DipoleError):
            call()


def test_argument_validation_without_a_gpu():
    """Bad arguments are rejected by the ABI before any CUDA call (return code + thread-local message)."""
    import dnp_native
    lib = dnp_native.load()
    rc = lib.dnp_field_grad_f32(None, -1, 6, None, 4, 3, 1e-5, None, 3, 0, None, 0, None)
    assert rc == -1 and b"negative size" in lib.dnp_last_error()
    rc = lib.dnp_field_grad_f32(None, 4, 6, None, 4, 3, 1e-5, None, 3, 0, None, 0, None)
    assert rc == -1 and b"null" in lib.dnp_last_error()
    one = ctypes.c_void_p(256)
    rc = lib.dnp_field_grad_f32(one, 4, 5, one, 4, 3, 1e-5, one, 3, 0, None, 0, None)
    assert rc == -1 and b"stride" in lib.dnp_last_error()
    rc = lib.dnp_field_sign_f32(one, 3, one, 3, None, 4, 0, None)
    assert rc == -1 and b"mode" in lib.dnp_last_error()
    rc = lib.dnp_points_propagate_f32(one, 10, 6, 11, 1e-6, 0, None, None, None, 0, None)
    assert rc == -1 and b"starting_point" in lib.dnp_last_error()
    assert lib.dnp_field_grad_f32(None, 0, 6, None, 0, 3, 1e-5, None, 3, 0, None, 0, None) == 0   # empty is OK


def test_window_kernel_reads_sources_through_uniform_registers():
    """The all-pairs kernels of dnp_upairs.cu are only fast while ptxas keeps the whole source index chain in the
    uniform datapath: every window read must be an LDCU.64 (constant bank -> uniform register) and every packed
    FP32 operation of the inner loop an FFMA2 / FMUL2 / FADD2.  Small changes to the kernel head have silently
    turned the LDCU.64 into vector LDC.64 before (see the note in dnp_upairs.cu); this guards the shipped binary."""
    import collections
    import shutil
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    sass = subprocess.check_output(["cuobjdump", "-sass", LIB], text=True)
    seen = 0
    for fn in re.split(r"\n\s+Function : ", sass)[1:]:
        name = fn.split("\n")[0]
        if "upairs_kernel" not in name and "upairs_tile_kernel" not in name:
            continue
        ins = [(int(m.group(1), 16), m.group(2).strip())
               for m in re.finditer(r"/\*([0-9a-f]{4,5})\*/\s+(.*?);", fn)]
        loops = []
        for addr, text in ins:
            m = re.search(r"BRA\S*\s+(?:\S+,\s+)?0x([0-9a-f]+)", text)
            if m and int(m.group(1), 16) < addr:
                loops.append((int(m.group(1), 16), addr))
        inner = min((l for l in loops if any("MUFU" in t for a, t in ins if l[0] <= a <= l[1])),
                    key=lambda l: l[1] - l[0])
        ops = collections.Counter(re.sub(r"^@!?U?P\d\s+", "", t).split()[0] for a, t in ins if inner[0] <= a <= inner[1])
        assert ops["LDCU.64"] >= 6 and ops["LDC.64"] == 0 and ops["LDC"] == 0 and ops["LDCU"] == 0, (name, dict(ops))
        assert ops["FFMA2.FTZ"] + ops["FMUL2.FTZ"] + ops["FADD2.FTZ"] >= 12 * 4 and ops["FFMA"] == 0, (name, dict(ops))
        assert ops["MUFU.RSQ"] > 0
        seen += 1
    assert seen == 7                    # field / potential x (8, 2 targets per thread shares, whole tiles) + ragged field tiles
