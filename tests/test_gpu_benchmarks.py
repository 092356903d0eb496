"""The larger circuits of the reference's hdl-benchmarks test-cases directory (north_star: "decrypted output wires bit-exact
... over all hdl-benchmarks test cases"): AES round core, chi-squared and Euclidean distance as gate netlists, the 32-bit
multiplier, the Euclidean distance and 5x5 matrix product as arithmetic netlists.  Yosys netlists go through the restated
preprocessor; inputs are the reference's CSV files; every wire that carries a value is compared with the plaintext
evaluation (Circuit::evaluate, src/circuit.rs:348-381)."""
import gzip
import os
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = os.path.join(G, "test-cases")


def _gz(*parts):
    with gzip.open(os.path.join(G, *parts), "rt") as f:
        return f.read()


@pytest.fixture(scope="module")
def bool_session():
    from helm_b200 import host as H
    s = H.Session("boolean_default", seed=31)
    yield s
    s.close()


def _run_gates(circ, sess, ins):
    want = circ.evaluate(ins)
    before = sess.stats()
    got, secs = sess.run(circ, ins, all_wires=True)
    after = sess.stats()
    bad = [w for w in want if bool(got[w]) != bool(want[w])]
    assert not bad, "wires differ from the plaintext evaluation: %s" % bad[:8]
    pbs = after["pbs_count"] - before["pbs_count"]
    assert pbs == circ.count(circ.PBS_GATES) + sum(1 for g in circ.gates() if g[1] == "mux")   # one pass, MUX = 2 PBS
    return got, secs, pbs


@pytest.mark.parametrize("netlist,case", [("aes_core_gates.v.gz", "aes.inputs.csv"),
                                          ("chi_squared_gates.v.gz", "chi_squared_bin_1.inputs.csv"),
                                          ("v2_euclidean_gates.v.gz", "v2-euclidean-distance-bin.inputs.csv")])
def test_gate_netlists_with_reference_inputs(bool_session, netlist, case):
    from helm_b200 import host as H
    c = H.Circuit(text=H.preprocess_yosys(_gz("yosys", netlist)))
    ins = H.read_input_wires(os.path.join(CASES, case))
    got, secs, pbs = _run_gates(c, bool_session, ins)
    w = c.level_widths()
    print("%s: %d gates, %d levels (max width %d), %d PBS, evaluate %.1f ms = %.0f PBS/s"
          % (netlist, c.count(c.GATES), len(w), max(w), pbs, 1e3 * secs[1], pbs / secs[1]))


def test_32_bit_multiplier(bool_session):
    from helm_b200 import host as H
    c = H.Circuit(text=_gz("netlists", "32-bit-mult-gates.v.gz"))
    ins = H.read_input_wires(os.path.join(CASES, "32-bit-mult.inputs.csv"))
    got, secs, pbs = _run_gates(c, bool_session, ins)
    a = sum(int(ins["G11[%d]" % i]) << i for i in range(32))
    b = sum(int(ins["G12[%d]" % i]) << i for i in range(32))
    assert sum(int(got["G14[%d]" % i]) << i for i in range(32)) == (a * b) & 0xFFFFFFFF
    print("32-bit-mult-gates: %d PBS in %d levels, evaluate %.1f ms" % (pbs, c.count(c.LEVELS), 1e3 * secs[1]))


def test_arithmetic_euclidean_distance_and_matrix_product():
    from helm_b200 import host as H
    s = H.Session("shortint_m2c2", seed=33)
    c = H.Circuit(text=H.preprocess_yosys(_gz("yosys", "v2-euclidean-distance.v.gz"), arith=True), is_arith=True)
    ins = H.read_input_wires(os.path.join(CASES, "v2-euclidean-distance.inputs.csv"), "u16")
    want = c.evaluate(ins, "u16")
    got, secs, cnt = s.run_arith(c, ins, "u16", all_wires=True)
    assert {k: got[k] for k in want} == want and want["dist"] == (57 - 13) ** 2 + (96 - 57) ** 2
    print("v2-euclidean-distance u16: %d bootstraps in %d batches, evaluate %.3f s" % (cnt["bootstraps"], cnt["batches"], secs[1]))
    m = H.Circuit(os.path.join(G, "netlists", "5x5_x_5x5-mmult-arith.v"), is_arith=True)
    rng = np.random.default_rng(6)
    A, B = rng.integers(0, 256, size=(5, 5)), rng.integers(0, 256, size=(5, 5))
    ins = {}
    for i in range(5):
        for j in range(5):
            ins["a_%d_%d" % (i, j)] = int(A[i, j])
            ins["b_%d_%d" % (i, j)] = int(B[i, j])
    want = m.evaluate(ins, "u8")
    got, secs, cnt = s.run_arith(m, ins, "u8")
    assert got and all(got[k] == want[k] for k in got)
    assert all(got["c_%d_%d" % (i, j)] == int((A @ B)[i, j]) % 256 for i in range(5) for j in range(5))
    print("5x5 matrix product u8: %d bootstraps in %d batches, evaluate %.3f s" % (cnt["bootstraps"], cnt["batches"], secs[1]))
    s.close()
