"""GPU parity of the text IO and wire format (util.py:46-69,
socket_server_para.py:145-166,191).  Byte / bit work: the parsed float32 rows and the exported text must be
IDENTICAL to the reference's (tests/golden/textio.npz) and to the oracle (Python's own float() / repr())."""
import inspect
import json
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _txt(g, key):
    return bytes(g[key]).decode()


def _same_bits(a, b):
    a, b = np.ascontiguousarray(a, dtype=np.float32), np.ascontiguousarray(b, dtype=np.float32)
    nan = np.isnan(a) & np.isnan(b)
    return a.shape == b.shape and np.array_equal(np.where(nan, 0, a.view(np.uint32)), np.where(nan, 0, b.view(np.uint32)))


def test_parse_matches_reference(golden):
    import dnp_util
    g = golden("textio")
    got = dnp_util.xyz2tensor(_txt(g, "text"))
    assert got.dtype == torch.float32 and not got.is_cuda
    assert _same_bits(got.numpy(), g["parsed_append"])
    assert _same_bits(dnp_util.xyz2tensor(_txt(g, "text_six"), append_normals=False).numpy(), g["parsed_six_noappend"])
    got3 = dnp_util.xyz2tensor(_txt(g, "text_three"), append_normals=False)
    assert got3.shape[1] == 3 and _same_bits(got3.numpy(), g["parsed_three_noappend"])
    assert dnp_util.xyz2tensor_cuda(_txt(g, "text")).is_cuda


def test_parse_errors_like_the_reference(golden):
    import dnp_util
    g = golden("textio")
    with pytest.raises(RuntimeError):                 # 3- and 6-column lines cannot be stacked (util.py:68)
        dnp_util.xyz2tensor(_txt(g, "text"), append_normals=False)
    with pytest.raises(RuntimeError):                 # nothing to stack
        dnp_util.xyz2tensor("\n\n1 2\n")
    with pytest.raises(ValueError):                   # float('') on a doubled space inside a 6-token line
        dnp_util.xyz2tensor("1  2 3 4 5\n")
    with pytest.raises(ValueError):
        dnp_util.xyz2tensor("1 2 x3\n")
    assert dnp_util.xyz2tensor("1 2 3").tolist() == [[1.0, 2.0, 3.0, 0.0, 0.0, 0.0]]       # no trailing newline


def test_parse_random_literals_against_python_float(dev):
    """Decimal -> double -> float32 against Python's own float() on 60 000 literals of every shape the parser
    accepts: random float32 reprs, %.{1..20}g renderings (long digit strings, rounding boundaries), exponents."""
    import dnp_util
    rng = np.random.default_rng(99)
    bits = rng.integers(0, 2 ** 32, 30000, dtype=np.uint64).astype(np.uint32)
    vals = bits.view(np.float32).astype(np.float64)
    vals = vals[np.isfinite(vals) & (np.abs(vals) < 1e38) & ((np.abs(vals) > 1e-37) | (vals == 0))]
    toks = [repr(float(v)) for v in vals]
    dbl = rng.normal(size=20000) * 10.0 ** rng.integers(-30, 30, 20000)
    toks += [("%." + str(int(p)) + "g") % v for v, p in zip(dbl, rng.integers(1, 21, 20000))]
    toks += [("%.17e") % v for v in dbl[:5000]]
    toks += ["%d.%019d" % (a, b) for a, b in zip(rng.integers(0, 1000, 3000), rng.integers(0, 10 ** 18, 3000))]
    toks = toks[:(len(toks) // 3) * 3]
    txt = "\n".join(" ".join(toks[i:i + 3]) for i in range(0, len(toks), 3))
    want = np.array([float(t) for t in toks]).astype(np.float32).reshape(-1, 3)
    got = dnp_util.xyz2tensor(txt, append_normals=False).numpy()
    assert _same_bits(got, want)


def test_export_matches_reference(golden, dev, tmp_path):
    import dnp_util
    g = golden("textio")
    pc = torch.from_numpy(g["export_pc"])
    want = bytes(g["export_text"])
    assert dnp_util.export_text(pc.transpose(0, 1).to(dev)) == want          # the [6, N] view the callers pass
    dest = tmp_path / "final_result.xyz"
    dnp_util.export_pc(pc.to(dev).transpose(0, 1), dest)                     # util.py:46 signature, CUDA cloud
    assert open(dest, "rb").read() == want
    dnp_util.export_pc(pc.transpose(0, 1).contiguous(), dest)                # CPU cloud, contiguous [6, N]
    assert open(dest, "rb").read() == want


def test_export_every_exponent_against_python_repr(orc, dev):
    """2 million random float32 bit patterns (all exponents, subnormals included) + the powers of two and ten."""
    import dnp_util
    rng = np.random.default_rng(3)
    bits = rng.integers(0, 2 ** 32, 2_000_000, dtype=np.uint64).astype(np.uint32)
    v = bits.view(np.float32).copy()
    v = v[~np.isnan(v)]
    v = np.concatenate([v, (2.0 ** np.arange(-149, 128)).astype(np.float32),
                        (10.0 ** np.arange(-45, 39)).astype(np.float32)])
    v = v[:(len(v) // 6) * 6].reshape(-1, 6)
    got = dnp_util.export_text(torch.from_numpy(v).to(dev).transpose(0, 1))
    want = orc.export_text(v).encode()
    if got != want:
        ga, wa = got.split(b"\n"), want.split(b"\n")
        bad = [(a, b) for a, b in zip(ga, wa) if a != b][:5]
        raise AssertionError(f"{len(ga)} vs {len(wa)} lines; first differences: {bad}")


def test_round_trip_at_full_size(dev):
    """1M points (BASELINE config 3's size): export -> parse returns the same float32 bits (size-independent
    property: repr round-trips), and the text has one line per point."""
    import dnp_native
    import dnp_util
    g = torch.Generator().manual_seed(8)
    pc = torch.randn(1 << 20, 6, generator=g).to(dev)
    text = dnp_util.export_text(pc.transpose(0, 1))
    assert text.count(b"\n") == (1 << 20) - 1
    back, mask, bad, lines = dnp_native.xyz_parse(text)
    assert bad < 0 and mask == 2 and lines == 1 << 20
    assert torch.equal(back.view(torch.int32), pc.view(torch.int32))


def test_export_to_file_streams_the_same_text(dev, tmp_path):
    """dnp_xyz_export_fd_f32 (what export_pc calls): the file holds exactly the bytes of export_text, for texts
    shorter than one 4 MB staging chunk, of a few chunks with a ragged tail, and of ~30 chunks; appending to an
    open descriptor works (the library only calls write(2)); a failed write is reported, not ignored."""
    import os
    import dnp_native
    import dnp_util
    g = torch.Generator().manual_seed(12)
    for n in (1, 7, 3000, 90001, 1 << 18):
        pc = (torch.randn(n, 6, generator=g) * 10.0 ** torch.randint(-8, 8, (n, 1), generator=g).float()).to(dev)
        want = dnp_util.export_text(pc.transpose(0, 1))
        dest = tmp_path / f"cloud_{n}.xyz"
        dnp_util.export_pc(pc.transpose(0, 1), dest)
        assert open(dest, "rb").read() == want, n
    with open(tmp_path / "two.xyz", "wb") as f:
        f.write(b"# header\n")
        f.flush()
        wrote = dnp_native.xyz_export_fd(pc[:5], f.fileno())
    body = open(tmp_path / "two.xyz", "rb").read()
    assert body == b"# header\n" + dnp_util.export_text(pc[:5].transpose(0, 1)) and wrote == len(body) - 9
    assert dnp_native.xyz_export_fd(pc[:0], 1) == 0
    fd = os.open(tmp_path / "readonly.xyz", os.O_CREAT | os.O_RDONLY)
    try:
        with pytest.raises(dnp_native.DipoleError, match="write failed"):
            dnp_native.xyz_export_fd(pc[:100], fd)
    finally:
        os.close(fd)
    # the library stays usable after the failure
    dnp_util.export_pc(pc[:100].transpose(0, 1), tmp_path / "after.xyz")
    assert open(tmp_path / "after.xyz", "rb").read() == dnp_util.export_text(pc[:100].transpose(0, 1))


def test_ok_xyz_rows(golden, dev):
    """The shipped demo cloud through the parser (first 2000 rows recorded from the reference)."""
    import dnp_util
    g = golden("textio")
    rows = g["ok_rows"]
    txt = "\n".join(" ".join(repr(float(x)) for x in r) for r in rows)
    assert _same_bits(dnp_util.xyz2tensor(txt).numpy(), rows)


def test_wire_format(dev):
    import dnp_util
    rng = np.random.default_rng(1)
    xyz = rng.normal(size=(5000, 3)) * 3.0
    rows = dnp_util.wire_decode(xyz.tobytes())                                # socket_server_para.py:166
    assert rows.shape == (5000, 6) and rows.dtype == torch.float32 and rows.is_cuda
    assert np.array_equal(rows.cpu().numpy()[:, :3], xyz.astype(np.float32))
    assert not rows[:, 3:].any()
    rows[:, 3:] = torch.randn(5000, 3, device=dev)
    reply = dnp_util.wire_encode(rows)                                        # socket_server_para.py:191
    assert reply == rows.cpu().numpy().astype(np.float64).tobytes()
    assert dnp_util.wire_decode(np.zeros((0, 3))).shape == (0, 6)


def test_signatures_match_the_reference():
    import dnp_util
    want = json.load(open(os.path.join(GOLDEN, "util_signatures.json")))
    for name in ("xyz2tensor", "export_pc", "npxyz2tensor"):
        got = [[p.name, None if p.default is inspect.Parameter.empty else repr(p.default)]
               for p in inspect.signature(getattr(dnp_util, name)).parameters.values()]
        assert got == want[name], name
