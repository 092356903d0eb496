"""north_star: "bit-exact decrypted outputs on every HELM benchmark netlist".  EVERY Yosys netlist of the reference's
hdl-benchmarks/netlists directory (ISCAS'85 c17..c7552 in gates mode and as `_lut`, `_lut_2`, `_lut_4`, `-lut-2-1` LUT netlists,
ISCAS'89 s-series, crc-32, the 16/32-bit multipliers, AES core, chi-squared, Euclidean distance) goes through the restated
preprocessor and levelizer and is evaluated on the GPU; every wire that carries a value is compared with the plaintext
evaluation (Circuit::evaluate, /root/reference/src/circuit.rs:348-381), on all-false inputs (the README invocation) and on
seeded random vectors (splitmix64 seed 0x48454C4D).  Sequential netlists run two cycles.

Parameter sets: gates -> tfhe boolean DEFAULT_PARAMETERS; LUT netlists with <= 2-input LUTs -> PARAM_MESSAGE_1_CARRY_1 (the
CLI's, /root/reference/src/bin/helm.rs:301); 3- and 4-input LUTs need msg*carry >= 2^arity -> PARAM_MESSAGE_2_CARRY_2
(SURVEY App. C-4); wider LUTs are rejected with an error."""
import numpy as np
import pytest

import corpus

pytestmark = pytest.mark.gpu
_SESS = {}


def _session(params):
    from helm_b200 import host as H
    if params not in _SESS:
        _SESS[params] = H.Session(params, seed=61)
    return _SESS[params]


def _vectors(c, count):
    import bench
    inputs = sorted(c.names(c.INPUTS))                    # declared inputs; the parser appends the DFF outputs
    r = bench.splitmix64(0x48454C4D, count * len(inputs)) & np.uint64(1)
    dffs = set(c.names(c.DFFS))
    yield None
    for v in range(count):
        yield {w: (False if w in dffs else bool(r[v * len(inputs) + i])) for i, w in enumerate(inputs)}


@pytest.mark.parametrize("name", corpus.names())
def test_corpus_netlist_every_wire(name):
    from helm_b200 import host as H
    if name in corpus.REFERENCE_PANICS:
        pytest.skip("reference panics on this netlist (constant DFF input): result undefined")
    c = H.Circuit(text=H.preprocess_yosys(corpus.text(name), exact_consts=True))
    gates = c.gates()
    arity = max([len(g[4]) for g in gates if g[1] == "lut"] + [0])
    lut = c.count(c.HAS_LUTS) != 0
    params = "boolean_default" if not lut else ("shortint_m1c1" if arity <= 2 else "shortint_m2c2")
    s = _session(params)
    if arity > 4:
        with pytest.raises(H.HostError, match="needs msg_mod\\*carry_mod"):
            s.run(c)
        return
    cycles = 2 if c.count(c.DFFS) > 1 else 1
    nvec = 16 if name in ("c880.v", "c7552.v") else (2 if c.count(c.GATES) < 4000 else 1)
    total = 0.0
    for ins in _vectors(c, nvec):
        want = c.evaluate(ins, cycles=cycles)
        got, secs = s.run(c, ins, cycles=cycles, all_wires=True)
        bad = [w for w in want if bool(got[w]) != bool(want[w])]
        assert not bad, "%s: wires differ from the plaintext evaluation: %s" % (name, bad[:8])
        total += secs[1]
    w = c.level_widths()
    print("%s under %s: %d gates, %d levels (max width %d), %d cycle(s), %d vectors, evaluate %.1f ms each"
          % (name, params, len(gates), len(w), max(w), cycles, nvec + 1, 1e3 * total / (nvec + 1)))
