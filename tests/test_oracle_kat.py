"""Known-answer tests of the CPU oracle (no GPU).  The reference pins this path at the DECRYPTED level only
(tests/gates_test.rs:9-108 truth tables; tests/circuit_test.rs); those tables are restated here.  Ciphertext-level
behaviour is pinned by algebraic properties and by the seeded golden digests under tests/golden/."""
import numpy as np
import pytest

from oracle import oracle as O


@pytest.fixture(scope="module")
def K(keys_factory):
    return keys_factory("toy32")


# ---- signed gadget decomposition
@pytest.mark.parametrize("bits,bl,lv", [(32, 6, 3), (32, 3, 4), (32, 7, 3), (32, 2, 8), (64, 18, 1), (64, 23, 1),
                                        (64, 3, 5), (64, 4, 4), (32, 8, 4), (64, 16, 4)])
def test_decompose_reconstructs_closest_representable(bits, bl, lv):
    rng = np.random.default_rng(bits + bl)
    xs = [0, 1, (1 << bits) - 1, 1 << (bits - 1), (1 << (bits - 1)) - 1] + [int(v) for v in rng.integers(0, 1 << 63, 200)]
    B, q = 1 << bl, 1 << bits
    for x in xs:
        x %= q
        d = O.decompose(bits, bl, lv, x)
        assert all(-B // 2 <= v <= B // 2 for v in d)
        rec = sum(v * (q >> (bl * (i + 1))) for i, v in enumerate(d)) % q
        shift = bits - bl * lv
        closest = (((x >> (shift - 1)) + 1) >> 1 << shift) % q if shift > 0 else x
        assert rec == closest, (x, d)


def test_decompose_known_answers():
    # hand-computed: B=2^6, l=3 on u32.  x = 0x80000000 -> state 0x20000 (18 bits); top digit res = 32 = B/2 with
    # nothing above it: carry = ((res-1)|state)&res = 0, so the digit stays +32 (balanced only when the next bit is set)
    assert O.decompose(32, 6, 3, 0x80000000) == [32, 0, 0]
    assert O.decompose(32, 6, 3, 0x80000000 + (1 << 26)) == [-31, 0, 0]
    assert O.decompose(32, 6, 3, 0) == [0, 0, 0]
    assert O.decompose(32, 6, 3, 1 << 26) == [1, 0, 0]
    assert O.decompose(32, 6, 3, 1 << 14) == [0, 0, 1]
    assert O.decompose(32, 6, 3, 1 << 13) == [0, 0, 1]          # rounds up to the closest representable
    assert O.decompose(32, 6, 3, (1 << 13) - 1) == [0, 0, 0]
    assert O.decompose(32, 6, 3, 0xFFFFFFFF) == [0, 0, 0]       # rounds to 2^32 = 0


# ---- gate truth tables (reference tests/gates_test.rs:9-108), both polynomial back-ends
@pytest.mark.parametrize("mode", [O.EXACT, O.FFT])
def test_gate_truth_tables(K, mode):
    f = {O.AND: lambda a, b: a & b, O.NAND: lambda a, b: 1 - (a & b), O.OR: lambda a, b: a | b,
         O.NOR: lambda a, b: 1 - (a | b), O.XOR: lambda a, b: a ^ b, O.XNOR: lambda a, b: 1 - (a ^ b)}
    seed = 0
    for op, fn in f.items():
        for a in (0, 1):
            for b in (0, 1):
                seed += 2
                out = K.bool_gate(op, K.enc_bool(a, seed), K.enc_bool(b, seed + 1), mode=mode)
                assert K.dec_bool(out) == bool(fn(a, b)), (op, a, b)
    for a in (0, 1):
        assert K.dec_bool(K.bool_gate(O.NOT, K.enc_bool(a, 90 + a))) == (not a)
        for b in (0, 1):
            for s in (0, 1):      # mux(then=a, else=b, sel=s): s ? a : b   (src/gates.rs:189-194, :265)
                out = K.bool_gate(O.MUX, K.enc_bool(a, 100), K.enc_bool(b, 101), K.enc_bool(s, 102), mode=mode)
                assert K.dec_bool(out) == bool(a if s else b)


def test_fft_backend_equals_exact_on_u32(K):
    """u32 torus: the folded f64 FFT rounds to the exact negacyclic product -> identical ciphertext words."""
    for seed in range(4):
        ct = K.enc_bool(seed & 1, 40 + seed)
        assert np.array_equal(K.pbs(ct, K.bool_tv(), mode=O.EXACT), K.pbs(ct, K.bool_tv(), mode=O.FFT))
    rng = np.random.default_rng(2)
    acc = rng.integers(0, 2 ** 32, size=(K.k + 1) * K.N, dtype=np.uint64).astype(np.uint32)
    assert np.array_equal(K.cmux(5, 777, acc, O.EXACT), K.cmux(5, 777, acc, O.FFT))


def test_cmux_selects_rotation(K):
    """CMUX semantics: acc' decrypts to acc * X^(a * s_i).  Checked through the GLWE phase of coefficient 0."""
    sk = K.lwe_sk()
    tv = K.bool_tv()
    for i in range(6):
        a = 2 * K.N - 1                       # rotate by X^(2N-1) = X^-1: coefficient 0 stays +1/8 either way
        out = K.cmux(i, a, tv, O.EXACT).reshape(K.k + 1, K.N)
        assert out.shape == (K.k + 1, K.N)
        a = K.N                               # X^N = -1: body flips sign iff s_i = 1
        out = K.cmux(i, a, tv, O.EXACT)
        lwe = np.zeros(K.kN + 1, dtype=np.uint32)
        o2 = out.reshape(K.k + 1, K.N)
        for j in range(K.k):
            lwe[j * K.N] = o2[j, 0]
            lwe[j * K.N + 1:(j + 1) * K.N] = (0 - o2[j, :0:-1].astype(np.int64)).astype(np.uint32)
        lwe[K.kN] = o2[K.k, 0]
        assert K.dec_bool(lwe, big=1) == (not sk[i])


def test_keyswitch_preserves_phase(K):
    for v in (0, 1):
        big = K.encrypt((1 if v else 7) << 29, 60 + v, big=1)
        small = K.keyswitch(big)
        d = (K.phase(small, 0) - K.phase(big, 1)) % (1 << 32)
        d = min(d, (1 << 32) - d)
        assert d < 1 << 24 and K.dec_bool(small) == bool(v)


def test_modswitch_range_and_rounding(K):
    ct = np.arange(K.n + 1, dtype=np.uint32) * np.uint32(0x01234567)
    ms = K.modswitch(ct)
    assert ms.max() <= 2 * K.N
    sh = 32 - 10                              # log2(2N) = 10 for N = 512
    assert all(int(m) == ((int(c) + (1 << (sh - 1))) >> sh) for m, c in zip(ms, ct))


@pytest.mark.parametrize("name", ["toy64", "toy64b"])
@pytest.mark.parametrize("mode", [O.EXACT, O.FFT])
def test_shortint_bivariate_lut(name, mode, keys_factory):
    """LUT mode, 2-input path (src/gates.rs:761-764): ct = 2*c0 + c1 -> KS -> PBS(table)."""
    Ks = keys_factory(name)
    mc = Ks.P.msg_mod * Ks.P.carry_mod
    for lut in (0x8, 0x6, 0x7):
        table = [(lut >> x) & 1 if x < 4 else 0 for x in range(mc)]
        for a in (0, 1):
            for b in (0, 1):
                lin = (Ks.enc_shortint(a, 11) * np.uint64(2) + Ks.enc_shortint(b, 12)).astype(np.uint64)
                out = Ks.shortint_lut(lin, table, mode=mode)
                assert Ks.dec_shortint(out) == (lut >> (2 * a + b)) & 1


def test_shortint_accumulator_layout(keys_factory):
    """generate_lookup_table (App. A.8): boxes of N/(msg*carry), first half-box negated and rotated to the end."""
    Ks = keys_factory("toy64")
    mc, N = 16, Ks.N
    table = list(range(16))
    tv = Ks.shortint_accumulator(table).reshape(Ks.k + 1, N)
    assert not tv[:Ks.k].any()
    box, d = N // mc, Ks.delta()
    body = tv[Ks.k]
    assert int(body[0]) == 0 and int(body[box // 2]) == d and int(body[box // 2 - 1]) == 0
    assert int(body[N - 1]) == 0                # -0
    assert int(body[N - box // 2 - 1]) == 15 * d
    tv1 = Ks.shortint_accumulator([1] * 16).reshape(Ks.k + 1, N)[Ks.k]
    assert int(tv1[N - 1]) == (1 << 64) - d and int(tv1[0]) == d
