"""Arithmetic mode on the GPU: decrypted results of radix-block programs run through
hb_eval_shortint_ops equal wrapping unsigned arithmetic -- the contract of tfhe-rs FheUintN that the reference's
ArithCircuit relies on (src/circuit.rs:1299-1454, goldens tests/circuit_test.rs:314-474)."""
import os
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
NET = os.path.join(HERE, "golden", "netlists")
CASES = os.path.join(HERE, "golden", "test-cases")

CELLS = """module m (a, b, c, y0, y1, y2, y3, y4, y5, y6, y7, y8, y9, y10, y11, y12, y13, y14, y15, y16);
  input [7:0] a;
  input [7:0] b;
  input [7:0] c;
  output [7:0] y0;
  add g0(a, b, y0);
  sub g1(a, b, y1);
  mult g2(a, b, y2);
  add g3(a, 77, y3);
  sub g4(b, 200, y4);
  mult g5(3, c, y5);
  mult g6(c, 100, y6);
  shl g7(a, 3, y7);
  shr g8(b, 5, y8);
  shl g9(c, 4, y9);
  mult g10(y0, y1, y10);
  add g11(y10, y2, y11);
  mult g12(a, a, y12);
  div g13(a, b, y13);
  div g14(c, 7, y14);
  shl g15(a, c, y15);
  shr g16(b, c, y16);
endmodule
"""


@pytest.fixture(scope="module")
def toy_session():
    from helm_b200 import host as H
    s = H.Session("toy64", seed=11)
    yield s
    s.close()


@pytest.mark.parametrize("ptxt_type", ["u8", "u16"])
def test_every_cell_type_matches_plain_arithmetic(toy_session, ptxt_type):
    from helm_b200 import host as H
    c = H.Circuit(text=CELLS, is_arith=True)
    rng = np.random.default_rng(5)
    bits = int(ptxt_type[1:])
    for trial in range(3):
        ins = {k: int(rng.integers(0, 1 << bits)) for k in "abc"}
        if trial == 0:
            ins = {"a": (1 << bits) - 1, "b": (1 << bits) - 1, "c": 0}          # every carry chain at full length
        if trial == 1:
            ins["b"] = 0                                                        # division by zero: all ones (tfhe-rs)
        want = c.evaluate(ins, ptxt_type)
        got, secs, cnt = toy_session.run_arith(c, ins, ptxt_type, all_wires=True)
        for w in want:
            assert got[w] == want[w], (ptxt_type, ins, w, got[w], want[w])
        assert cnt["bootstraps"] > 0 and cnt["batches"] < cnt["bootstraps"] // 4   # batched, not one launch per block


def test_block_ops_through_the_c_abi(keys_factory, engine_factory):
    """LINEAR and LUT block operations with multi-input linear parts, checked block by block."""
    from helm_b200 import engine as E
    K = keys_factory("toy64")
    eng = engine_factory("toy64")
    rng = np.random.default_rng(9)
    n = 24
    vals = rng.integers(0, 4, size=(n, 3))
    w = eng.wires(3 * n + 2 * n)
    cts = np.stack([K.enc_shortint(int(vals[g, j]), 500 + 3 * g + j) for g in range(n) for j in range(3)])
    w.write(np.arange(3 * n, dtype=np.uint32), cts)
    ops = np.zeros(2 * n, dtype=E.SOP_DTYPE)
    table = 0
    for i in range(16):
        table |= ((i * 7 + 3) % 16) << (4 * i)
    for g in range(n):
        # LUT over a + b + c + 2  (three sources, constant)
        ops[g]["kind"] = E.SOP_LUT; ops[g]["n_in"] = 3; ops[g]["coef"][:3] = 1; ops[g]["cst"] = 2
        ops[g]["in"][:3] = [3 * g, 3 * g + 1, 3 * g + 2]; ops[g]["out"] = 3 * n + g; ops[g]["table"] = table
        # LINEAR 4*a - b + 3
        o = ops[n + g]
        o["kind"] = E.SOP_LINEAR; o["n_in"] = 2; o["coef"][:2] = [4, -1]; o["cst"] = 3
        o["in"][:2] = [3 * g, 3 * g + 1]; o["out"] = 4 * n + g
    eng.eval_shortint_ops(w, ops)
    eng.sync()
    got = w.read(np.arange(3 * n, 5 * n, dtype=np.uint32))
    for g in range(n):
        a, b, c = (int(v) for v in vals[g])
        assert K.dec_shortint_full(got[g]) == ((a + b + c + 2) * 7 + 3) % 16
        assert K.dec_shortint_full(got[n + g]) == (4 * a - b + 3) % 16
    w.close()


@pytest.mark.parametrize("i", [1, 2, 3, 4])
def test_chi_squared_goldens_u32(i):
    """BASELINE.json configs[3]: chi_squared_arith.v at u32 under PARAM_MESSAGE_2_CARRY_2 against the reference's
    golden CSVs (tests/circuit_test.rs:314-474)."""
    from helm_b200 import host as H
    s = _m2c2()
    c = H.Circuit(os.path.join(NET, "chi_squared_arith.v"), is_arith=True)
    ins = H.read_input_wires(os.path.join(CASES, "chi_squared_arith_%d.inputs.csv" % i), "u32")
    want = H.read_input_wires(os.path.join(CASES, "chi_squared_arith_%d.outputs.csv" % i), "u32")
    got, secs, cnt = s.run_arith(c, ins, "u32")
    assert {k: got[k] for k in want} == want
    print("chi_squared u32 case %d: %d bootstraps in %d batches, evaluate %.3f s" % (i, cnt["bootstraps"], cnt["batches"], secs[1]))


_M2C2 = []


def _m2c2():
    from helm_b200 import host as H
    if not _M2C2:
        _M2C2.append(H.Session("shortint_m2c2", seed=21))
    return _M2C2[0]


WIDE = """module m (a, b, y0, y1, y2, y3, y4, y5);
  input [7:0] a;
  input [7:0] b;
  output [7:0] y0;
  add g0(a, b, y0);
  sub g1(a, b, y1);
  mult g2(a, b, y2);
  shr g3(y2, 9, y3);
  mult g4(y1, 1000003, y4);
  add g5(y4, 4000000007, y5);
endmodule
"""


@pytest.mark.parametrize("ptxt_type", ["u64", "u128"])
def test_wide_integer_types(toy_session, ptxt_type):
    """FheUint64 / FheUint128 (reference src/gates.rs:386-600 dispatches on the plaintext type): 32 and 64 radix blocks."""
    from helm_b200 import host as H
    c = H.Circuit(text=WIDE, is_arith=True)
    bits = int(ptxt_type[1:])
    ins = {"a": (1 << bits) - 3, "b": (1 << (bits - 1)) + 12345}
    want = c.evaluate(ins, ptxt_type)
    assert want["y2"] == (ins["a"] * ins["b"]) % (1 << bits)                 # the plaintext evaluator itself
    got, secs, cnt = toy_session.run_arith(c, ins, ptxt_type, all_wires=True)
    for w in want:
        assert got[w] == want[w], (ptxt_type, w, got[w], want[w])
    print("%s: %d bootstraps in %d batches, evaluate %.3f s" % (ptxt_type, cnt["bootstraps"], cnt["batches"], secs[1]))
