"""READY-latched outputs of sequential circuits (reference EvalCircuit::{init_ready, evaluate_ready}: gates mode
src/circuit.rs:482-504, LUT mode :1002-1030, arithmetic :1195-1297; driven by the cycle loop of src/bin/helm.rs:256-270)
and the DFF level's pre-level semantics (every DFF latches the value its D wire had BEFORE the level, as the reference's
clone-under-read-lock / write-under-write-lock loop gives, src/circuit.rs:532-540) -- chained DFFs form a shift register.

Expected values come from the plaintext evaluator run cycle by cycle: valid[o] starts false / 0 and becomes the output's
value in every cycle whose READY wire is true / 1."""
import os

import pytest

pytestmark = pytest.mark.gpu
NET = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "netlists")


def _dff_init(c, zero):
    return {g[3]: zero for g in c.gates() if g[1] == "dff"}


def _latched_plain(c, ins, cycles, ptxt, outs, zero):
    valid = {o: zero for o in outs}
    for cy in range(1, cycles + 1):
        v = c.evaluate(ins, ptxt, cycles=cy)
        if v["READY"]:
            valid = {o: v[o] for o in outs}
    return valid


@pytest.mark.parametrize("netlist,params", [("ready_gates.v", "boolean_default"), ("ready_lut.v", "shortint_m1c1"),
                                            ("ready_lut.v", "toy64b")])
def test_ready_latching_boolean_modes(netlist, params):
    """2-bit counter raising READY in cycle 3 only: after 1-2 cycles the latched outputs are still the initial false, after
    3 and 4 they hold the cycle-3 values (not the current ones)."""
    from helm_b200 import host as H
    c = H.Circuit(os.path.join(NET, netlist))
    s = H.Session(params, seed=41)
    outs = ["READY", "y", "z"]
    for a, b in ((True, True), (False, True), (True, False)):
        ins = {"a": a, "b": b}
        ins.update(_dff_init(c, False))
        for cycles in (2, 3, 4):
            want = _latched_plain(c, ins, cycles, "bool", outs, False)
            got, _ = s.run(c, ins, cycles=cycles)
            assert {o: bool(got[o]) for o in outs} == {o: bool(want[o]) for o in outs}, (a, b, cycles)
        # the current (unlatched) wires after 4 cycles differ from the latched ones: y is a ^ q1
        cur, _ = s.run(c, ins, cycles=4, all_wires=True)
        plain = c.evaluate(ins, cycles=4)
        assert all(bool(cur[w]) == bool(plain[w]) for w in plain)
    s.close()


def test_ready_latching_arithmetic():
    """arithmetic mode: total accumulates x every cycle, READY = (cycle - 1) >> 1; valid = total*READY + valid*(1-READY)."""
    from helm_b200 import host as H
    c = H.Circuit(os.path.join(NET, "ready_arith.v"), is_arith=True)
    s = H.Session("toy64", seed=43)
    for x in (5, 77):
        ins = {"x": x}
        ins.update(_dff_init(c, 0))
        for cycles, want in ((2, 0), (3, 3 * x), (4, (4 * x) % 256)):
            plain = _latched_plain(c, ins, cycles, "u8", ["total"], 0)
            assert plain["total"] == want
            got, _, _ = s.run_arith(c, ins, "u8", cycles=cycles)
            assert got["total"] == want, (x, cycles, got)
    s.close()


@pytest.mark.parametrize("params", ["boolean_default", "toy32"])
def test_chained_dffs_shift_register(params):
    """q1 <= !d, q2 <= q1, q3 <= q2 in ONE DFF level: each DFF must see the pre-level value of its source (ADVICE r1:
    a single in-place copy launch could mix old and new words)."""
    from helm_b200 import host as H
    c = H.Circuit(os.path.join(NET, "shift_register.v"))
    s = H.Session(params, seed=45)
    for d in (False, True):
        ins = {"d": d}
        ins.update(_dff_init(c, False))
        for cycles in (1, 2, 3, 4):
            want = c.evaluate(ins, cycles=cycles)
            got, _ = s.run(c, ins, cycles=cycles, all_wires=True)
            assert {w: bool(got[w]) for w in want} == {w: bool(v) for w, v in want.items()}, (d, cycles)
    s.close()
