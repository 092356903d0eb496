"""GPU parity of the two primitives under every gate (reference src/gates.rs:255-274 -> tfhe-rs
programmable bootstrap + keyswitch) against the CPU oracle, through the C ABI."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

U32_SETS = ["toy32", "boolean_default"]


@pytest.mark.parametrize("name", U32_SETS + ["toy64", "toy64b"])
def test_keyswitch_bit_exact(name, keys_factory, engine_factory):
    K, eng = keys_factory(name), engine_factory(name)
    rng = np.random.default_rng(1)
    cnt = 37      # ragged: not a multiple of the 16-gate tile
    info = np.iinfo(K.dtype)
    cts = rng.integers(0, info.max, size=(cnt, K.kN + 1), dtype=K.dtype, endpoint=True)
    got = eng.keyswitch_batch(cts)
    for i in range(cnt):
        assert np.array_equal(got[i], K.keyswitch(cts[i])), "keyswitch row %d differs" % i


@pytest.mark.parametrize("name", U32_SETS)
def test_pbs_bit_exact_u32(name, keys_factory, engine_factory, oracle_mod):
    """u32 torus: the f64 FFT external product rounds to the exact integer negacyclic product,
    so the GPU ciphertext must be bit-identical to the oracle's exact-integer back-end."""
    K, eng = keys_factory(name), engine_factory(name)
    cnt = 5 if name == "boolean_default" else 9     # odd: exercises the half-filled G=2 CTA
    cts = np.stack([K.enc_bool(i & 1, 100 + i) for i in range(cnt)])
    tv = K.bool_tv().reshape(K.k + 1, K.N)[K.k]
    got = eng.pbs_batch(cts, tv[None, :])
    for i in range(cnt):
        want = K.pbs(cts[i], K.bool_tv(), mode=oracle_mod.EXACT)
        assert np.array_equal(got[i], want), "pbs row %d differs in %d words" % (i, np.count_nonzero(got[i] != want))
        assert K.dec_bool(got[i], big=1) == bool(i & 1)


def test_pbs_arbitrary_test_vector(keys_factory, engine_factory, oracle_mod):
    K, eng = keys_factory("toy32"), engine_factory("toy32")
    rng = np.random.default_rng(5)
    cnt = 4
    cts = rng.integers(0, 2 ** 32, size=(cnt, K.n + 1), dtype=np.uint64).astype(np.uint32)
    tvb = rng.integers(0, 2 ** 32, size=(cnt, K.N), dtype=np.uint64).astype(np.uint32)
    got = eng.pbs_batch(cts, tvb)
    for i in range(cnt):
        tv = np.zeros((K.k + 1, K.N), dtype=np.uint32)
        tv[K.k] = tvb[i]
        want = K.pbs(cts[i], tv.reshape(-1), mode=oracle_mod.EXACT)
        assert np.array_equal(got[i], want)


@pytest.mark.parametrize("name", ["toy64", "toy64b"])
def test_pbs_u64_close_to_exact(name, keys_factory, engine_factory, oracle_mod):
    """u64 torus: the f64 FFT cannot be bit-exact (53-bit mantissa), exactly as in tfhe-rs.  The output phase must
    agree with the exact-integer oracle to well below the scheme noise, and with the oracle's own f64-FFT
    back-end to the same order."""
    K, eng = keys_factory(name), engine_factory(name)
    rng = np.random.default_rng(9)
    cnt = 3
    cts = rng.integers(0, 2 ** 63, size=(cnt, K.n + 1), dtype=np.uint64) * np.uint64(2)
    mc = K.P.msg_mod * K.P.carry_mod
    tv = K.shortint_accumulator(list(range(mc))).reshape(K.k + 1, K.N)
    got = eng.pbs_batch(cts, tv[K.k][None, :])
    for i in range(cnt):
        want = K.pbs(cts[i], tv.reshape(-1), mode=oracle_mod.EXACT)
        d = (got[i].astype(np.int64) - want.astype(np.int64))          # wrapping difference per word
        worst = int(np.abs(d).max())
        assert worst < 1 << 40, "word error 2^%.1f (torus 2^%.1f)" % (np.log2(max(worst, 1)), np.log2(max(worst, 1)) - 64)
        ph = (K.phase(got[i], 1) - K.phase(want, 1)) % (1 << 64)
        assert min(ph, (1 << 64) - ph) < K.delta() >> 12
