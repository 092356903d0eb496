"""Committed known-answer vectors (tests/golden/oracle/*.npz, written by scripts/gen_oracle_goldens.py).
CPU part: the oracle still reproduces them (pins the restatement against silent changes).
GPU part: the CUDA engine reproduces them through the C ABI -- bit for bit on the u32 torus, and on the u64 torus the
integer keyswitch bit for bit and the f64-FFT bootstraps to the same decryption with phases within delta/64."""
import os
import numpy as np
import pytest

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "oracle")


def _load(name):
    return np.load(os.path.join(G, name + ".npz"))


def test_decomposition_goldens(oracle_mod):
    z = _load("decomposition")
    for key in z.files:
        if not key.startswith("x_"):
            continue
        bits, bl, lv = (int(v) for v in key.split("_")[1:])
        want = z["d_%d_%d_%d" % (bits, bl, lv)]
        for x, d in zip(z[key], want):
            assert oracle_mod.decompose(bits, bl, lv, int(x)) == [int(v) for v in d]
            rec = sum(int(v) << (bits - bl * (i + 1)) for i, v in enumerate(d)) % (1 << bits)      # closest representable
            err = (rec - int(x)) % (1 << bits)
            err = min(err, (1 << bits) - err)
            assert err <= 1 << (bits - bl * lv - 1)


@pytest.mark.parametrize("name", ["toy32", "boolean_default"])
def test_oracle_reproduces_boolean_goldens(name, keys_factory, oracle_mod):
    z = _load(name)
    K = keys_factory(name, int(z["key_seed"]))
    ops, cts = z["ops"], z["inputs"]
    ng = len(ops) if name == "toy32" else 2                       # the full-size set costs ~1 s per gate on the CPU
    for g in range(ng):
        got = K.bool_gate(int(ops[g]), cts[3 * g], cts[3 * g + 1], cts[3 * g + 2], mode=oracle_mod.EXACT)
        assert np.array_equal(got, z["gate_outputs"][g])
        assert K.dec_bool(got) == bool({0: lambda x, y, s: x & y, 1: lambda x, y, s: 1 - (x & y), 2: lambda x, y, s: x | y,
                                        3: lambda x, y, s: 1 - (x | y), 4: lambda x, y, s: x ^ y, 5: lambda x, y, s: 1 - (x ^ y),
                                        6: lambda x, y, s: x if s else y}[int(ops[g])](*(int(v) for v in z["bits"][g])))
    lin = K.bool_linear(int(ops[0]), cts[0], cts[1])
    assert np.array_equal(lin, z["linear"]) and np.array_equal(K.modswitch(lin), z["modswitch"])
    assert np.array_equal(K.cmux(int(z["cmux_step"]), int(z["cmux_a"]), z["cmux_in"], mode=oracle_mod.EXACT), z["cmux_out"])
    assert np.array_equal(K.keyswitch(z["pbs"]), z["keyswitch"])
    if name == "toy32":
        assert np.array_equal(K.pbs(lin, K.bool_tv(), mode=oracle_mod.EXACT), z["pbs"])
        assert np.array_equal(K.pbs(lin, K.bool_tv(), mode=oracle_mod.FFT), z["pbs"])     # f64 FFT is exact on the u32 torus


@pytest.mark.parametrize("name", ["toy64b", "toy64", "toy64m"])
def test_oracle_reproduces_shortint_goldens(name, keys_factory, oracle_mod):
    z = _load(name)
    K = keys_factory(name, int(z["key_seed"]))
    table = [int(v) for v in z["table"]]
    for i in range(2):
        assert np.array_equal(K.keyswitch(z["inputs"][i]), z["keyswitch"][i])
        out = K.shortint_lut(z["inputs"][i], table, mode=oracle_mod.FFT)
        assert K.dec_shortint(out) == table[int(z["values"][i])] == K.dec_shortint(z["lut_outputs_fft"][i])
    if "lut_outputs_exact" in z:        # multi-bit blind rotation: the exact-integer back-end is deterministic word for word, and the
        ex = K.shortint_lut(z["inputs"][0], table, mode=oracle_mod.EXACT)          # f64-FFT back-end stays inside the output noise of it
        assert np.array_equal(ex, z["lut_outputs_exact"][0])
        for i in range(len(z["inputs"])):
            d = (int(K.phase(z["lut_outputs_exact"][i], 1)) - int(K.phase(z["lut_outputs_fft"][i], 1))) % (1 << 64)
            assert min(d, (1 << 64) - d) < K.delta() // 64
            assert K.dec_shortint(z["lut_outputs_exact"][i]) == table[int(z["values"][i])]


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["toy32", "boolean_default"])
def test_engine_reproduces_boolean_goldens(name, engine_factory):
    from helm_b200 import engine as E
    z = _load(name)
    eng = engine_factory(name, int(z["key_seed"]))
    ops, cts = z["ops"], z["inputs"]
    ng = len(ops)
    w = eng.wires(4 * ng)
    w.write_rows(0, cts)
    ar = np.arange(ng)
    hb_ops = np.array([E.AND, E.NAND, E.OR, E.NOR, E.XOR, E.XNOR, E.MUX], dtype=np.uint8)[ops]
    eng.eval_level(w, E.make_gates(hb_ops, 3 * ar, 3 * ar + 1, 3 * ng + ar, in2=3 * ar + 2))
    eng.sync()
    got = np.zeros((ng, eng.ct_words), dtype=eng.dtype)
    w.read_rows(3 * ng, got)
    assert np.array_equal(got, z["gate_outputs"])
    w.close()
    tv = np.zeros(eng.P.N, dtype=eng.dtype); tv[:] = 1 << (eng.P.torus_bits - 3)
    assert np.array_equal(eng.pbs_batch(z["linear"][None, :], tv[None, :])[0], z["pbs"])
    assert np.array_equal(eng.keyswitch_batch(z["pbs"][None, :])[0], z["keyswitch"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["toy64b", "toy64", "toy64m"])
def test_engine_reproduces_shortint_goldens(name, keys_factory, engine_factory):
    from helm_b200 import engine as E
    z = _load(name)
    K, eng = keys_factory(name, int(z["key_seed"])), engine_factory(name, int(z["key_seed"]))
    cts, table = z["inputs"], [int(v) for v in z["table"]]
    n = len(cts)
    assert np.array_equal(eng.keyswitch_batch(cts), z["keyswitch"])                       # integer path: bit-exact
    w = eng.wires(2 * n)
    w.write_rows(0, cts)
    ops = np.zeros(n, dtype=E.SOP_DTYPE)
    packed = sum((table[i] & 15) << (4 * i) for i in range(len(table)))
    for i in range(n):
        ops[i]["kind"] = E.SOP_LUT; ops[i]["n_in"] = 1; ops[i]["coef"][0] = 1; ops[i]["in"][0] = i
        ops[i]["out"] = n + i; ops[i]["table"] = packed
    eng.eval_shortint_ops(w, ops)
    eng.sync()
    got = np.zeros((n, eng.ct_words), dtype=eng.dtype)
    w.read_rows(n, got)
    d = K.delta()
    for i in range(n):
        assert K.dec_shortint(got[i]) == table[int(z["values"][i])]
        diff = (int(K.phase(got[i], 1)) - int(K.phase(z["lut_outputs_fft"][i], 1))) % (1 << 64)
        diff = min(diff, (1 << 64) - diff)
        assert diff < d // 64
    w.close()
