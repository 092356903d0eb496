"""Multi-bit (grouping factor 3) blind rotation -- the reference's ACTUAL arithmetic-mode parameter set,
PARAM_MULTI_BIT_MESSAGE_2_CARRY_2_GROUP_3_KS_PBS (/root/reference/src/bin/helm.rs:81-86; SURVEY 8f item 2, VERDICT r1 missing #2):
n = 888 -> 296 sequential external products, 2^3 GGSWs per group combined with the monomial spectra of the switched mask sums.

Checked against the CPU oracle's restatement (exact-integer and f64-FFT back-ends, oracle/tfhe_oracle.cpp::multibit_step_t) on the
same keys and inputs: decryptions equal the table, phases agree inside the scheme's own output noise; output-noise variance on the
real set; the reference's chi-squared golden CSVs at u32 through arithmetic mode (tests/circuit_test.rs:314-474 uses u16)."""
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _signed(x):
    return np.asarray(x, dtype=np.uint64).astype(np.int64).astype(np.float64) / 2.0 ** 64


def _lut_level(K, eng, W, seed):
    """W LUT ops v = 4*a + b -> table[v] through hb_eval_shortint_ops; returns (messages, inputs, tables, outputs)."""
    from helm_b200 import engine as EE
    rng = np.random.default_rng(seed)
    msgs = rng.integers(0, 4, size=(W, 2))
    cts = K.encrypt_batch(msgs.reshape(-1).astype(np.uint64) * np.uint64(K.delta()), 777000 + seed, big=1)
    w = eng.wires(3 * W)
    w.write_rows(0, cts)
    ops = np.zeros(W, dtype=EE.SOP_DTYPE)
    ops["kind"] = EE.SOP_LUT
    ops["n_in"] = 2
    ops["coef"][:, 0] = 4
    ops["coef"][:, 1] = 1
    ops["in"][:, 0] = 2 * np.arange(W)
    ops["in"][:, 1] = 2 * np.arange(W) + 1
    ops["out"] = 2 * W + np.arange(W)
    tabs = rng.integers(0, 16, size=(W, 16))
    ops["table"] = [sum(int(t) << (4 * i) for i, t in enumerate(row)) for row in tabs]
    eng.eval_shortint_ops(w, ops)
    eng.sync()
    out = np.zeros((W, eng.ct_words), dtype=eng.dtype)
    w.read_rows(2 * W, out)
    w.close()
    return msgs, cts, tabs, out


@pytest.mark.parametrize("W", [40, 333])
def test_multibit_lut_against_oracle(W, keys_factory, engine_factory, oracle_mod):
    """narrow (one gate per CTA) and wide instances; decrypt = table entry; phase vs the oracle's exact-integer back-end."""
    O = oracle_mod
    K, eng = keys_factory("toy64m"), engine_factory("toy64m")
    msgs, cts, tabs, out = _lut_level(K, eng, W, 5)
    idx = 4 * msgs[:, 0] + msgs[:, 1]
    want = tabs[np.arange(W), idx]
    got = np.array([K.dec_shortint_full(c) for c in out])
    assert np.array_equal(got, want)
    ns = min(W, 8)

    def ref(g):
        lin = (cts[2 * g] * K.dtype(4) + cts[2 * g + 1]).astype(K.dtype)
        return K.phase(K.shortint_lut(lin, [int(t) for t in tabs[g]], mode=O.EXACT), 1)
    with ThreadPoolExecutor(8) as ex:
        ph_ref = np.array(list(ex.map(ref, range(ns))), dtype=np.uint64)
    d = np.abs(_signed(K.phase_batch(out[:ns], big=1) - ph_ref))
    assert d.max() < K.delta() / 2.0 ** 64 / 64, "phase differs from the exact oracle by 2^%.1f" % np.log2(d.max())


def test_multibit_real_parameters_noise(keys_factory, engine_factory, oracle_mod):
    """The real set: 512 LUT bootstraps decrypt correctly, output-noise variance below the bound (n/g steps, each an external
    product with a sum of 2^g GGSWs: Var ~ (n/g) [2^g l (k+1) N (B^2+2)/12 sigma^2 + (1 + kN/2) / (12 B^2l)]) and in line with the
    oracle's f64-FFT path on a sample; the keyswitch has n + 1 = 889 output columns (tensor-core path only)."""
    O = oracle_mod
    name = "shortint_multibit_m2c2_g3"
    K, eng = keys_factory(name), engine_factory(name)
    W = 512
    msgs, cts, tabs, out = _lut_level(K, eng, W, 9)
    idx = 4 * msgs[:, 0] + msgs[:, 1]
    want = tabs[np.arange(W), idx].astype(np.uint64)
    err = _signed(K.phase_batch(out, big=1) - want * np.uint64(K.delta()))
    half = K.delta() / 2.0 ** 65
    assert np.abs(err).max() < half / 4
    P = K.params
    steps = P["n"] // P["grouping"]
    B = 2.0 ** P["pbs_base_log"]
    bound = steps * ((1 << P["grouping"]) * (P["k"] + 1) * P["N"] * (B * B + 2) / 12 * P["glwe_std"] ** 2
                     + (1 + P["k"] * P["N"] / 2) / (12 * B * B))
    var = float(err.var())
    assert var < 1.5 * bound, "variance 2^%.2f vs bound 2^%.2f" % (np.log2(var), np.log2(bound))

    def ref(g):
        lin = (cts[2 * g] * K.dtype(4) + cts[2 * g + 1]).astype(K.dtype)
        return K.phase(K.shortint_lut(lin, [int(t) for t in tabs[g]], mode=O.FFT), 1)
    with ThreadPoolExecutor(16) as ex:
        ph = np.array(list(ex.map(ref, range(48))), dtype=np.uint64)
    var_cpu = float(_signed(ph - want[:48] * np.uint64(K.delta())).var())
    assert 0.35 * var_cpu < var < 3.0 * var_cpu
    print("%s: %d LUT bootstraps, sigma 2^%.2f (bound 2^%.2f, oracle f64-FFT 2^%.2f), half step 2^%.1f"
          % (name, W, 0.5 * np.log2(var), 0.5 * np.log2(bound), 0.5 * np.log2(var_cpu), np.log2(half)))


_S = {}


def _sess(params):
    from helm_b200 import host as H
    if params not in _S:
        _S[params] = H.Session(params, seed=91)
    return _S[params]


@pytest.mark.parametrize("case", [1, 2, 3, 4])
def test_chi_squared_u32_under_the_reference_parameter_set(case):
    """configs[3] under the parameter set the reference really builds arithmetic mode on; classic M2C2 latency beside it."""
    from helm_b200 import host as H
    c = H.Circuit(os.path.join(G, "netlists", "chi_squared_arith.v"), is_arith=True)
    ins = H.read_input_wires(os.path.join(G, "test-cases", "chi_squared_arith_%d.inputs.csv" % case), "u32")
    want = H.read_input_wires(os.path.join(G, "test-cases", "chi_squared_arith_%d.outputs.csv" % case), "u32")
    res = {}
    for params in ("shortint_multibit_m2c2_g3", "shortint_m2c2"):
        s = _sess(params)
        got, secs, cnt = s.run_arith(c, ins, "u32")
        assert {k: got[k] for k in want} == want, params
        res[params] = (secs[1], cnt)
    print("chi_squared u32 case %d: multi-bit G3 %.3f s, classic M2C2 %.3f s (%d bootstraps in %d batches)"
          % (case, res["shortint_multibit_m2c2_g3"][0], res["shortint_m2c2"][0], res["shortint_m2c2"][1]["bootstraps"],
             res["shortint_m2c2"][1]["batches"]))
