"""Host-side logic without a GPU: the reference's parser / circuit tests restated against the C++ host library
(tests/verilog_parser_test.rs:4-141, tests/circuit_test.rs:17-45) plus the netlist statistics of SURVEY.md App. B."""
import os

import pytest

from helm_b200 import host as H

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
NET = os.path.join(G, "netlists")
CASES = os.path.join(G, "test-cases")
C = H.Circuit


def test_header_symbols_exported():
    import re
    src = open(os.path.join(os.path.dirname(G), "..", "include", "helm_host.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    syms = sorted(set(re.findall(r"\b(hh_[a-z0-9_]+)\s*\(", src)))
    L = H.lib()
    for s in syms:
        assert hasattr(L, s), s
    assert sorted(H.EXPORTS) == syms


def test_parse_two_bit_adder():                       # verilog_parser_test.rs:4-13
    c = C(os.path.join(NET, "2-bit-adder.v"))
    assert c.count(C.GATES) == 10 and c.count(C.WIRES) == 10 and c.count(C.INPUTS) == 5
    assert c.count(C.HAS_LUTS) == 0 and c.count(C.HAS_ARITH) == 0


def test_input_wires_parser():                        # verilog_parser_test.rs:15-27, 60-75
    w = H.read_input_wires(os.path.join(CASES, "2-bit-adder.inputs.csv"))
    assert w == {"a[0]": True, "a[1]": False, "b[0]": False, "b[1]": True, "cin": False}
    c = C(os.path.join(NET, "2-bit-adder.v"))
    assert set(w) == {g for g in c.evaluate(w) if "[" in g and g[0] in "ab"} | {"cin"}


def test_arithmetic_parser_and_int_types():           # verilog_parser_test.rs:29-47, 77-118
    c = C(os.path.join(NET, "chi_squared_arith.v"), is_arith=True)
    assert c.count(C.HAS_ARITH) == 1 and c.count(C.GATES) == 14 and c.count(C.INPUTS) == 3 and c.count(C.OUTPUTS) == 4
    for t in ("u8", "u16", "u32", "u64", "u128"):
        assert H.read_input_wires(os.path.join(CASES, "chi_squared_arith_1.inputs.csv"), t) == {"N0": 2, "N1": 7, "N2": 9}


def test_invalid_netlist_mixes_luts_and_arith():      # verilog_parser_test.rs:49-58 (#[should_panic])
    with pytest.raises(H.HostError, match="mix LUTs with arithmetic"):
        C(os.path.join(NET, "invalid.v"), is_arith=True)


def test_bool_array_csv_hex_expansion():              # verilog_parser_test.rs:120-141
    w = H.read_input_wires(os.path.join(CASES, "bool_array.input.csv"))
    assert [w["in1[%d]" % i] for i in range(4)] == [True, False, True, True]
    assert [w["in2[%d]" % i] for i in range(4)] == [False, True, False, True]              # 0xA LSB first
    assert [w["in3[%d]" % i] for i in range(6)] == [True, True, False, True, False, False]  # 0xB zero padded
    assert len(w) == 14


def test_two_bit_adder_plaintext():                   # circuit_test.rs:17-45
    c = C(os.path.join(NET, "2-bit-adder.v"))
    v = c.evaluate({k: True for k in ("a[0]", "a[1]", "b[0]", "b[1]", "cin")})
    assert len(v) == 15
    assert v["sum[0]"] and v["sum[1]"] and v["cout"] and not v["i0"] and not v["i1"]
    assert [g[2] for g in c.gates() if g[0] == 1] == ["a1", "a3", "x1", "x3"]      # level order is by name


@pytest.mark.parametrize("name,levels,widths,pbs", [
    ("s27.v", 6, [4, 3, 1, 1, 2, 3], 7),
    ("c17.v", 4, None, 6),
    ("16-bit-mult-gates.v", 31, None, 684),
    ("8-bit-adder-lut-3-1.v", 8, None, 15),
    ("8-bit-adder-lut-2-1.v", 14, None, 34),
])
def test_levelization_statistics(name, levels, widths, pbs):
    c = C(os.path.join(NET, name))
    assert c.count(C.LEVELS) == levels and c.count(C.PBS_GATES) == pbs
    if widths:
        assert c.level_widths() == widths


def test_s27_all_false_and_sequential():              # BASELINE config 1: no inputs => all false, G17 = true
    c = C(os.path.join(NET, "s27.v"))
    assert c.evaluate()["G17"] is True
    v3 = c.evaluate(cycles=3)
    assert isinstance(v3["G17"], bool) and c.count(C.DFFS) == 3


def test_lut_adders_compute_54():                     # 19 + 35 (8-bit-adder.inputs.csv); the stale outputs csv says 4
    ins = H.read_input_wires(os.path.join(CASES, "8-bit-adder.inputs.csv"))
    for name in ("8-bit-adder-lut-2-1.v", "8-bit-adder-lut-3-1.v"):
        c = C(os.path.join(NET, name))
        v = c.evaluate(ins)
        assert sum(int(v["G14[%d]" % i]) << i for i in range(8)) == 54, name


@pytest.mark.parametrize("i", [1, 2, 3, 4])
def test_chi_squared_goldens_plain(i):                # circuit_test.rs:314-474 golden CSVs (valid at u16 and u32)
    c = C(os.path.join(NET, "chi_squared_arith.v"), is_arith=True)
    for t in ("u16", "u32"):
        ins = H.read_input_wires(os.path.join(CASES, "chi_squared_arith_%d.inputs.csv" % i), t)
        want = H.read_input_wires(os.path.join(CASES, "chi_squared_arith_%d.outputs.csv" % i), t)
        got = c.evaluate(ins, t)
        assert {k: got[k] for k in want} == want


def test_numeric_wire_names_quirk():
    txt = "module m(a, b, y);\n input a, b;\n output y;\n and g2(a, b, 10);\n not g1(10, y);\nendmodule\n"
    ok = C(text=txt)
    assert ok.level_widths() == [1, 1] and ok.evaluate({"a": True, "b": True})["y"] is False
    quirk = C(text=txt, strict_reference_levels=True)      # consumer sorts first by name: the reference puts both gates in level 1 (a race)
    assert quirk.level_widths() == [2]
    with pytest.raises(H.HostError, match="never become ready"):
        C(text="module m(a, y);\n input a;\n output y;\n and g1(a, zz, y);\nendmodule\n")


def test_yosys_preprocessor_matches_checked_in_c17():  # hdl-benchmarks/src/lib.rs:319-391, 608-810
    txt = H.preprocess_yosys(open(os.path.join(os.path.dirname(NET), "yosys", "c17.v")).read())
    mine = C(text=txt)
    ref = C(os.path.join(NET, "c17.v"))                # produced by the reference's own preprocessor
    strip = lambda s: s.strip("_")                     # the checked-in file predates keeping Yosys' underscores
    a = sorted((g[1], strip(g[3]), tuple(strip(w) for w in g[4])) for g in mine.gates())
    b = sorted((g[1], strip(g[3]), tuple(strip(w) for w in g[4])) for g in ref.gates())
    assert a == b and mine.level_widths() == ref.level_widths()
    bits = {"G1": True, "G2": False, "G3": True, "G4": True, "G5": False}
    va, vb = mine.evaluate(bits), ref.evaluate(bits)
    assert va["G16"] == vb["G16"] and va["G17"] == vb["G17"]


@pytest.mark.parametrize("name,levels,maxw,pbs_gates,mux", [("c880", 26, 51, 262, 0), ("c7552", 31, 180, 902, 135)])
def test_iscas85_shapes(name, levels, maxw, pbs_gates, mux):
    c = C(text=H.preprocess_yosys(open(os.path.join(os.path.dirname(NET), "yosys", name + ".v")).read()))
    assert c.count(C.LEVELS) == levels and max(c.level_widths()) == maxw and c.count(C.PBS_GATES) == pbs_gates
    assert sum(1 for g in c.gates() if g[1] == "mux") == mux
    c.evaluate()                                        # all-false inputs evaluate without missing wires


def test_client_rng_is_chacha20_known_answer():
    """client randomness (keys, masks, noise) must come from a CSPRNG.  The generator is the ChaCha20 block
    function; RFC 8439 section 2.3.2 test vector (key 00..1f, counter 1, nonce 00:00:00:09 00:00:00:4a 00:00:00:00)."""
    import ctypes
    import numpy as np
    key = np.frombuffer(bytes(range(32)), dtype="<u4").copy()
    w = np.array([1, 0x09000000, 0x4a000000, 0], dtype=np.uint32)
    out = np.zeros(16, dtype=np.uint32)
    H.lib().hh_chacha20_block(key.ctypes.data, w.ctypes.data, out.ctypes.data)
    want = [0xe4e7f110, 0x15593bd1, 0x1fdd0f50, 0xc47120a3, 0xc7f4d1c7, 0x0368c033, 0x9aaa2204, 0x4e6cd4c3,
            0x466482d2, 0x09aa9f07, 0x05d7c214, 0xa2028bd9, 0xd19c12b5, 0xb94e16de, 0xe883d0cb, 0x4e3c50a2]
    assert [int(v) for v in out] == want


def test_sequential_fixtures_plaintext():
    """The sequential fixtures of tests/test_gpu_ready.py on the plaintext evaluator: a chain of DFFs is a shift register
    (each DFF latches the pre-level value of its D wire), and the 2-bit counter raises READY in cycle 3 only."""
    c = C(os.path.join(NET, "shift_register.v"))
    assert c.count(C.DFFS) == 3 and c.level_widths() == [1, 3]
    ins = {"d": False, "q1": False, "q2": False, "q3": False}
    assert [c.evaluate(ins, cycles=k)["q3"] for k in (1, 2, 3, 4)] == [False, False, True, True]
    for name in ("ready_gates.v", "ready_lut.v"):
        r = C(os.path.join(NET, name))
        ins = {"a": True, "b": False, "q0": False, "q1": False, "q2": False}
        assert [r.evaluate(ins, cycles=k)["READY"] for k in (1, 2, 3, 4)] == [False, False, True, False]
        assert r.evaluate(ins, cycles=3)["y"] is False and r.evaluate(ins, cycles=4)["y"] is False
    a = C(os.path.join(NET, "ready_arith.v"), is_arith=True)
    ins = {"x": 9, "cnt": 0, "acc": 0}
    assert [(a.evaluate(ins, "u8", cycles=k)["READY"], a.evaluate(ins, "u8", cycles=k)["total"]) for k in (1, 2, 3, 4)] == \
        [(0, 9), (0, 18), (1, 27), (1, 36)]


def test_corpus_preprocesses_and_levelizes():
    """Every Yosys netlist of the reference's corpus (tests/golden/yosys) goes through the restated preprocessor, levelizer and
    plaintext evaluator; SURVEY App. B shapes for the ISCAS'85 gate netlists (HELM convention: NOT is a level)."""
    import corpus
    names = corpus.names()
    assert len(names) >= 68
    shapes = {"c432.v": (30, 148), "c499.v": (11, 200), "c880.v": (26, 262), "c1355.v": (11, 198), "c1908.v": (17, 157),
              "c6288.v": (73, 1593), "16-bit-mult-gates.v": (31, 684)}
    for name in names:
        if name in corpus.REFERENCE_PANICS:
            continue
        c = C(text=H.preprocess_yosys(corpus.text(name), exact_consts=True))
        assert c.count(C.LEVELS) >= 1 and c.evaluate()
        if name in shapes:
            assert (c.count(C.LEVELS), c.count(C.PBS_GATES)) == shapes[name], name


def test_not_folding_batch_counts():
    """NOT / BUF contracted out of the encrypted schedule (Circuit::folded_batches): fewer sequential batches than reference levels
    where a NOT sits on the critical path; never more."""
    import corpus
    want = {"c432.v": (30, 26), "c7552.v": (31, 30), "c2670.v": (19, 18), "c880.v": (26, 26), "16-bit-mult-gates.v": (31, 31)}
    for name, (levels, batches) in want.items():
        c = C(text=H.preprocess_yosys(corpus.text(name), exact_consts=True))
        assert (c.count(C.LEVELS), c.count(C.BATCHES)) == (levels, batches), name
    s27 = C(os.path.join(NET, "s27.v"))
    assert s27.count(C.BATCHES) <= s27.count(C.LEVELS)
