"""CPU tests of the text IO row: the oracle's restatement of util.xyz2tensor /
util.export_pc against outputs of the reference itself (tests/golden/textio.npz)."""
import numpy as np


def _txt(g, key):
    return bytes(g[key]).decode()


def test_oracle_parse_matches_reference(golden, orc):
    g = golden("textio")
    got = orc.xyz2tensor(_txt(g, "text"), append_normals=True)
    assert got.dtype == np.float32 and np.array_equal(got, g["parsed_append"], equal_nan=True)
    assert np.array_equal(orc.xyz2tensor(_txt(g, "text_six"), False), g["parsed_six_noappend"], equal_nan=True)
    got3 = orc.xyz2tensor(_txt(g, "text_three"), False)
    assert got3.shape[1] == 3 and np.array_equal(got3, g["parsed_three_noappend"], equal_nan=True)


def test_oracle_export_matches_reference(golden, orc):
    g = golden("textio")
    assert orc.export_text(g["export_pc"]).encode() == bytes(g["export_text"])


def test_export_then_parse_round_trips(golden, orc):
    """repr is the shortest text that reads back as the same double, hence the same float32."""
    g = golden("textio")
    pc = g["export_pc"]
    keep = np.isfinite(pc).all(axis=1)
    back = orc.xyz2tensor(orc.export_text(pc[keep]))
    assert np.array_equal(back.view(np.uint32), pc[keep].view(np.uint32))


def test_reference_style_loops_rate(orc, capsys):
    """Not a speed assertion: records what the reference's per-line Python loops cost on this host (the figure
    DESIGN.md section 3.11 quotes next to the kernels)."""
    import time
    rng = np.random.default_rng(0)
    pc = rng.normal(size=(20000, 6)).astype(np.float32)
    t0 = time.perf_counter()
    txt = orc.export_text(pc)
    t1 = time.perf_counter()
    back = orc.xyz2tensor(txt)
    t2 = time.perf_counter()
    assert np.array_equal(back, pc)
    with capsys.disabled():
        print(f"\n[reference-style text IO on this host] export {1e6 * (t1 - t0) / pc.size:.2f} us/value, "
              f"parse {1e6 * (t2 - t1) / pc.size:.2f} us/value, {len(txt) / pc.size:.1f} bytes/value")
