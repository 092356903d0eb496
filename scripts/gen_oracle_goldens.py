#!/usr/bin/env python
"""Generates tests/golden/oracle/*.npz: seeded known-answer vectors of the CPU oracle (oracle/, exact-integer mode) for
the primitives of the hot path -- signed decomposition, modulus switch, one CMUX (external product), one full PBS, one
keyswitch, boolean gates and one shortint LUT -- per parameter set ("golden vectors to generate").

Why: the reference holds no ciphertext-level vectors (encryption is randomised, tfhe-rs is not buildable here), so these
pin (1) the oracle against silent changes of itself and (2) the CUDA engine against the oracle at fixed keys, through files
that travel to the GPU box.  Keys are regenerated from the seed by the oracle's deterministic generator; only inputs and
expected outputs are stored.  Run here: python scripts/gen_oracle_goldens.py"""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle as O          # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "oracle")
KEY_SEED = 3


def boolean_set(name, n_gates):
    K = O.Keys(name, seed=KEY_SEED)
    rng = np.random.default_rng(2024)
    bits = rng.integers(0, 2, size=(n_gates, 3))
    ops = np.arange(n_gates) % 7                      # AND NAND OR NOR XOR XNOR MUX
    cts = K.enc_bool_batch(bits.reshape(-1), 424242)
    outs = np.stack([K.bool_gate(int(ops[g]), cts[3 * g], cts[3 * g + 1], cts[3 * g + 2], mode=O.EXACT) for g in range(n_gates)])
    # primitives on the first gate's linear part
    lin = K.bool_linear(int(ops[0]), cts[0], cts[1])
    ms = K.modswitch(lin)
    pbs = K.pbs(lin, K.bool_tv(), mode=O.EXACT)
    ks = K.keyswitch(pbs)
    N, k = K.P.N, K.P.k
    acc = np.zeros((k + 1) * N, dtype=K.dtype)
    acc[k * N:] = rng.integers(0, 1 << K.bits, size=N, dtype=np.uint64).astype(K.dtype)
    cm = K.cmux(0, int(ms[0]) if int(ms[0]) else 5, acc.copy(), mode=O.EXACT)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), key_seed=KEY_SEED, bits=bits, ops=ops, inputs=cts, gate_outputs=outs,
                        linear=lin, modswitch=ms, pbs=pbs, keyswitch=ks, cmux_step=0, cmux_a=int(ms[0]) if int(ms[0]) else 5,
                        cmux_in=acc, cmux_out=cm)


def shortint_set(name, n_luts):
    K = O.Keys(name, seed=KEY_SEED)
    rng = np.random.default_rng(77)
    mc = K.P.msg_mod * K.P.carry_mod
    vals = rng.integers(0, K.P.msg_mod, size=n_luts)
    table = [(3 * i + 1) % K.P.msg_mod for i in range(mc)]
    cts = np.stack([K.enc_shortint(int(v), 9000 + i) for i, v in enumerate(vals)])
    outs = np.stack([K.shortint_lut(cts[i], table, mode=O.FFT) for i in range(n_luts)])
    extra = {}
    if K.P.grouping > 1:                              # exact-integer outputs as well: the two back-ends must stay within the output noise
        extra["lut_outputs_exact"] = np.stack([K.shortint_lut(cts[i], table, mode=O.EXACT) for i in range(n_luts)])
    np.savez_compressed(os.path.join(OUT, name + ".npz"), key_seed=KEY_SEED, values=vals, table=np.array(table), inputs=cts,
                        lut_outputs_fft=outs, keyswitch=np.stack([K.keyswitch(c) for c in cts]), **extra)


def decomposition():
    rng = np.random.default_rng(1)
    rows = {}
    for bits, bl, lv in ((32, 6, 3), (32, 3, 4), (32, 7, 3), (32, 2, 8), (64, 18, 1), (64, 23, 1), (64, 3, 5), (64, 4, 4)):
        xs = rng.integers(0, 1 << 63, size=256, dtype=np.uint64) * 2 + rng.integers(0, 2, size=256, dtype=np.uint64)
        if bits == 32:
            xs = xs & np.uint64(0xffffffff)
        edge = np.array([0, 1, (1 << bits) - 1, 1 << (bits - 1), (1 << (bits - bl * lv)) - 1, 1 << (bits - bl * lv - 1)], dtype=np.uint64)
        xs = np.concatenate([edge, xs])
        rows["x_%d_%d_%d" % (bits, bl, lv)] = xs
        rows["d_%d_%d_%d" % (bits, bl, lv)] = np.stack([O.decompose(bits, bl, lv, int(x)) for x in xs])
    np.savez_compressed(os.path.join(OUT, "decomposition.npz"), **rows)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    decomposition()
    boolean_set("toy32", 14)
    boolean_set("boolean_default", 7)
    shortint_set("toy64b", 6)
    shortint_set("toy64", 4)
    shortint_set("toy64m", 4)                         # multi-bit (grouping 3) blind rotation, oracle/tfhe_oracle.cpp::multibit_step_t
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))
