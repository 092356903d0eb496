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


@pytest.mark.parametrize("name", ["toy32", "boolean_default", "concrete_gpu", "toy64", "toy64b", "shortint_m1c1", "shortint_m2c2"])
def test_keyswitch_wide_batch_tensor_core_bit_exact(name, keys_factory, engine_factory):
    """Batches from 96 rows on run the int8 tensor-core keyswitch (ks_mma_kernel.cuh: byte planes of the key, 4 for the u32
    torus and 8 for u64, any ks_level): word for word equal to the oracle, ragged against the 64-row tile, random full-range
    inputs (every digit value and every carry chain of the signed decomposition occurs)."""
    K, eng = keys_factory(name), engine_factory(name)
    rng = np.random.default_rng(2)
    cnt = 203
    info = np.iinfo(K.dtype)
    cts = rng.integers(0, info.max, size=(cnt, K.kN + 1), dtype=K.dtype, endpoint=True)
    cts[0, :] = 0; cts[1, :] = info.max; cts[2, :] = info.max // 2 + 1          # edge rows
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
def test_pbs_u64_phase_parity(name, keys_factory, engine_factory, oracle_mod):
    """u64 torus: the f64 FFT cannot be bit-exact (53-bit mantissa), exactly as in tfhe-rs, and a single flipped
    rounding in one decomposition digit changes every later ciphertext word -- so words are not comparable, phases
    are.  (1) With <= 1 non-zero mask element there is no feedback: words agree with the exact-integer oracle to
    the f64 rounding level.  (2) On real encryptions the decrypted phase equals the looked-up value and sits within
    the scheme's noise of the exact oracle's phase."""
    K, eng = keys_factory(name), engine_factory(name)
    mc = K.P.msg_mod * K.P.carry_mod
    tv = K.shortint_accumulator(list(range(mc))).reshape(K.k + 1, K.N)
    ct = np.zeros(K.n + 1, dtype=np.uint64)
    ct[3] = np.uint64(0x4000000000000000); ct[K.n] = np.uint64(12345 << 50)
    got = eng.pbs_batch(ct[None, :], tv[K.k][None, :])[0]
    want = K.pbs(ct, tv.reshape(-1), mode=oracle_mod.EXACT)
    assert int(np.abs((got - want).astype(np.int64)).max()) < 1 << 44
    small = np.stack([K.encrypt(m * K.delta(), 300 + m, big=0) for m in range(mc // 2)])
    got = eng.pbs_batch(small, tv[K.k][None, :])
    for m in range(mc // 2):
        want = K.pbs(small[m], tv.reshape(-1), mode=oracle_mod.EXACT)
        d = (K.phase(got[m], 1) - K.phase(want, 1)) % (1 << 64)
        assert min(d, (1 << 64) - d) < K.delta() >> 6
        ideal = (K.phase(got[m], 1) - m * K.delta()) % (1 << 64)
        assert min(ideal, (1 << 64) - ideal) < K.delta() >> 4
