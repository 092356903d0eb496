"""The reference's own GPU parameter set and its GPU tests (row a12): `concrete_gpu` = n 512, k 1, N 1024, PBS l 3 / B 2^7,
KS l 8 / B 2^2, u32 torus (/root/reference/src/bin/helm.rs:141-146, the concrete-core CudaEngine path of
src/circuit.rs:650-934) and the 16- / 32-bit multipliers of tests/circuit_test.rs:96-264.  Everything is compared at the
CIPHERTEXT level with the CPU oracle on the same keys and input ciphertexts (u32 torus: the f64 transform is exact, so
ORC_FFT == ORC_EXACT word for word -- asserted below on a level of every gate type), and at the decrypted level with the
plaintext netlist."""
import gzip
import os

import numpy as np
import pytest

from circuit_util import run_both

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_concrete_gpu_level_all_gate_types_bit_exact(keys_factory, engine_factory, oracle_mod):
    from helm_b200 import engine as E
    O = oracle_mod
    K, eng = keys_factory("concrete_gpu"), engine_factory("concrete_gpu")
    ops = [E.AND, E.NAND, E.OR, E.NOR, E.XOR, E.XNOR, E.MUX, E.NOT, E.COPY]
    ng = len(ops)
    bits = np.random.default_rng(5).integers(0, 2, size=(ng, 3))
    cts = K.enc_bool_batch(bits.reshape(-1), 31000)
    w = eng.wires(4 * ng)
    w.write(np.arange(3 * ng, dtype=np.uint32), cts)
    ar = np.arange(ng)
    eng.eval_level(w, E.make_gates(ops, 3 * ar, 3 * ar + 1, 3 * ng + ar, in2=3 * ar + 2))
    eng.sync()
    got = w.read((3 * ng + ar).astype(np.uint32))
    for g, op in enumerate(ops):
        exact = K.bool_gate(op, cts[3 * g], cts[3 * g + 1], cts[3 * g + 2], mode=O.EXACT)
        fft = K.bool_gate(op, cts[3 * g], cts[3 * g + 1], cts[3 * g + 2], mode=O.FFT)
        assert np.array_equal(exact, fft), "oracle: f64-FFT and exact-integer back-ends differ on the u32 torus"
        assert np.array_equal(got[g], exact), "gate %d (op %d): ciphertext differs from the exact oracle" % (g, op)
    w.close()


def _mult_case(path, csv, nbits, keys_factory, engine_factory, O):
    from helm_b200 import engine as E, host as H
    K, eng = keys_factory("concrete_gpu"), engine_factory("concrete_gpu")
    text = gzip.open(path, "rt").read() if path.endswith(".gz") else open(path).read()
    c = H.Circuit(text=text)
    ins = H.read_input_wires(csv)
    before = eng.stats()
    idx, gpu, cpu = run_both(c, K, eng, E, O, ins, 52000, O.FFT)
    pbs = eng.stats()["pbs_count"] - before["pbs_count"]
    assert np.array_equal(gpu, cpu), "wire ciphertexts differ from the oracle: rows %s" % np.nonzero((gpu != cpu).any(axis=1))[0][:8]
    want = c.evaluate(ins)
    dec = K.dec_bool_batch(gpu)
    bad = [w for w in want if w in idx and bool(dec[idx[w]]) != bool(want[w])]
    assert not bad, "decrypted wires differ from the plaintext netlist: %s" % bad[:8]
    a = sum(int(ins["G11[%d]" % i]) << i for i in range(nbits))
    b = sum(int(ins["G12[%d]" % i]) << i for i in range(nbits))
    assert sum(int(dec[idx["G14[%d]" % i]]) << i for i in range(nbits)) == (a * b) % (1 << nbits)
    return pbs


def test_encrypted_16_bit_multiplier_concrete_gpu(keys_factory, engine_factory, oracle_mod):   # circuit_test.rs:96-180
    pbs = _mult_case(os.path.join(G, "netlists", "16-bit-mult-gates.v"), os.path.join(G, "test-cases", "16-bit-mult.inputs.csv"),
                     16, keys_factory, engine_factory, oracle_mod)
    assert pbs == 684


def test_encrypted_32_bit_multiplier_concrete_gpu(keys_factory, engine_factory, oracle_mod):   # circuit_test.rs:182-264
    pbs = _mult_case(os.path.join(G, "netlists", "32-bit-mult-gates.v.gz"), os.path.join(G, "test-cases", "32-bit-mult.inputs.csv"),
                     32, keys_factory, engine_factory, oracle_mod)
    assert pbs == 2962
