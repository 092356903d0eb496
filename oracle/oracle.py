"""ctypes binding of the CPU oracle (oracle/tfhe_oracle.h).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  Never imported by helm_b200/.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

EXACT, FFT = 0, 1
AND, NAND, OR, NOR, XOR, XNOR, MUX, NOT, COPY, LUT = range(10)


class Params(C.Structure):
    _fields_ = [("n", C.c_int32), ("k", C.c_int32), ("N", C.c_int32),
                ("pbs_base_log", C.c_int32), ("pbs_level", C.c_int32),
                ("ks_base_log", C.c_int32), ("ks_level", C.c_int32),
                ("torus_bits", C.c_int32), ("order", C.c_int32),
                ("msg_mod", C.c_int32), ("carry_mod", C.c_int32), ("grouping", C.c_int32),
                ("lwe_std", C.c_double), ("glwe_std", C.c_double)]


# Parameter sets of SURVEY.md section 8 (values recalled from tfhe-rs 0.4.x; run-time data, not code)
PARAM_SETS = {
    # tfhe::boolean DEFAULT_PARAMETERS (gen_keys(), reference src/bin/helm.rs:241)
    "boolean_default": dict(n=722, k=2, N=512, pbs_base_log=6, pbs_level=3, ks_base_log=3, ks_level=4,
                            torus_bits=32, order=0, msg_mod=0, carry_mod=0,
                            lwe_std=1.3071021089943935e-5, glwe_std=4.990272175010415e-8),
    # PARAM_MESSAGE_1_CARRY_1_KS_PBS (reference src/bin/helm.rs:301)
    "shortint_m1c1": dict(n=684, k=3, N=512, pbs_base_log=18, pbs_level=1, ks_base_log=3, ks_level=4,
                          torus_bits=64, order=1, msg_mod=2, carry_mod=2,
                          lwe_std=2.0399883307556e-5, glwe_std=3.45253423096e-13),
    # PARAM_MESSAGE_2_CARRY_2_KS_PBS
    "shortint_m2c2": dict(n=742, k=1, N=2048, pbs_base_log=23, pbs_level=1, ks_base_log=3, ks_level=5,
                          torus_bits=64, order=1, msg_mod=4, carry_mod=4,
                          lwe_std=7.069849454709433e-6, glwe_std=2.94036403378e-16),
    # concrete-core GPU path of the reference (src/bin/helm.rs:141-146)
    "concrete_gpu": dict(n=512, k=1, N=1024, pbs_base_log=7, pbs_level=3, ks_base_log=2, ks_level=8,
                         torus_bits=32, order=0, msg_mod=0, carry_mod=0,
                         lwe_std=2.0 ** -25, glwe_std=2.0 ** -25),
    # PARAM_MULTI_BIT_MESSAGE_2_CARRY_2_GROUP_3_KS_PBS: the reference's arithmetic-mode set (src/bin/helm.rs:81-86); multi-bit
    # blind rotation, grouping factor 3 (values as recalled in SURVEY.md section 8 / App. A.1, medium confidence)
    "shortint_multibit_m2c2_g3": dict(n=888, k=1, N=2048, pbs_base_log=21, pbs_level=1, ks_base_log=4, ks_level=4,
                                      torus_bits=64, order=1, msg_mod=4, carry_mod=4, grouping=3,
                                      lwe_std=6.125031601933181e-7, glwe_std=2.9403601535432533e-16),
    "toy64m": dict(n=36, k=1, N=2048, pbs_base_log=21, pbs_level=1, ks_base_log=4, ks_level=4,
                   torus_bits=64, order=1, msg_mod=4, carry_mod=4, grouping=3, lwe_std=2.0 ** -24, glwe_std=2.0 ** -52),
    # small set for fast CPU tests (NOT secure; same code paths, k=2 / l=3 like boolean_default)
    "toy32": dict(n=48, k=2, N=512, pbs_base_log=6, pbs_level=3, ks_base_log=3, ks_level=4,
                  torus_bits=32, order=0, msg_mod=0, carry_mod=0, lwe_std=2.0 ** -20, glwe_std=2.0 ** -28),
    "toy64": dict(n=40, k=1, N=2048, pbs_base_log=23, pbs_level=1, ks_base_log=3, ks_level=5,
                  torus_bits=64, order=1, msg_mod=4, carry_mod=4, lwe_std=2.0 ** -24, glwe_std=2.0 ** -52),
    "toy64b": dict(n=36, k=3, N=512, pbs_base_log=18, pbs_level=1, ks_base_log=3, ks_level=4,
                   torus_bits=64, order=1, msg_mod=2, carry_mod=2, lwe_std=2.0 ** -24, glwe_std=2.0 ** -42),
}


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        vp, u64, u32p = C.c_void_p, C.c_uint64, C.POINTER(C.c_uint32)
        L.orc_keygen.restype = vp
        L.orc_keygen.argtypes = [C.POINTER(Params), u64]
        L.orc_keys_free.argtypes = [vp]
        L.orc_bsk.restype = vp
        L.orc_bsk.argtypes = [vp, C.POINTER(C.c_size_t)]
        L.orc_ksk.restype = vp
        L.orc_ksk.argtypes = [vp, C.POINTER(C.c_size_t)]
        L.orc_lwe_sk.restype = vp
        L.orc_lwe_sk.argtypes = [vp]
        L.orc_glwe_sk.restype = vp
        L.orc_glwe_sk.argtypes = [vp]
        L.orc_encrypt.argtypes = [vp, C.c_int, u64, u64, vp]
        L.orc_trivial.argtypes = [vp, C.c_int, u64, vp]
        L.orc_phase.restype = u64
        L.orc_phase.argtypes = [vp, C.c_int, vp]
        L.orc_encrypt_batch.argtypes = [vp, C.c_int, vp, u64, C.c_uint32, vp]
        L.orc_phase_batch.argtypes = [vp, C.c_int, vp, C.c_uint32, vp]
        L.orc_decompose.argtypes = [C.c_int, C.c_int, C.c_int, u64, C.POINTER(C.c_int64)]
        L.orc_keyswitch.argtypes = [vp, vp, vp]
        L.orc_pbs.argtypes = [vp, vp, vp, vp, C.c_int]
        L.orc_cmux.argtypes = [vp, C.c_int, C.c_uint32, vp, C.c_int]
        L.orc_modswitch.argtypes = [vp, vp, vp]
        L.orc_bool_gate.argtypes = [vp, C.c_int, vp, vp, vp, vp, C.c_int]
        L.orc_bool_linear.argtypes = [vp, C.c_int, vp, vp, vp]
        L.orc_shortint_accumulator.argtypes = [vp, vp, vp]
        L.orc_shortint_lut.argtypes = [vp, vp, vp, vp, C.c_int]
        L.orc_set_threads.restype = C.c_int
        L.orc_set_threads.argtypes = [C.c_int]
        L.orc_bool_level.restype = C.c_int
        L.orc_bool_level.argtypes = [vp, vp, vp, vp, vp, vp, vp, C.c_uint32, C.c_int]
        _LIB = L
    return _LIB


def set_threads(n=0):
    """OpenMP team of the batched calls (0: every online core, whatever OMP_NUM_THREADS says); returns the team size."""
    return int(lib().orc_set_threads(n))


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class Keys:
    """Client + server key material for one parameter set (seeded, deterministic)."""

    def __init__(self, params, seed=3):
        if isinstance(params, str):
            params = PARAM_SETS[params]
        self.params = dict(params)
        self.params.setdefault("grouping", 0)
        self.P = Params(**self.params)
        self.L = lib()
        self.h = self.L.orc_keygen(C.byref(self.P), seed)
        self.dtype = np.uint32 if self.P.torus_bits == 32 else np.uint64
        self.n, self.k, self.N = self.P.n, self.P.k, self.P.N
        self.kN = self.k * self.N
        self.bits = self.P.torus_bits

    def __del__(self):
        try:
            self.L.orc_keys_free(self.h)
        except Exception:
            pass

    def _view(self, fn):
        sz = C.c_size_t()
        ptr = fn(self.h, C.byref(sz))
        buf = (C.c_uint8 * sz.value).from_address(ptr)
        return np.frombuffer(buf, dtype=self.dtype)

    def bsk(self):
        g = self.P.grouping
        lead = (self.n // g) * (1 << g) if g > 1 else self.n       # multi-bit: [group][pattern] flattened
        return self._view(self.L.orc_bsk).reshape(lead, self.P.pbs_level, self.k + 1, self.k + 1, self.N)

    def ksk(self):
        return self._view(self.L.orc_ksk).reshape(self.kN, self.P.ks_level, self.n + 1)

    def lwe_sk(self):
        return np.frombuffer((C.c_uint8 * self.n).from_address(self.L.orc_lwe_sk(self.h)), dtype=np.uint8)

    def glwe_sk(self):
        return np.frombuffer((C.c_uint8 * self.kN).from_address(self.L.orc_glwe_sk(self.h)), dtype=np.uint8)

    def dim(self, big):
        return self.kN if big else self.n

    def encrypt(self, pt, seed, big=0):
        ct = np.zeros(self.dim(big) + 1, dtype=self.dtype)
        self.L.orc_encrypt(self.h, big, int(pt) & ((1 << self.bits) - 1), seed, _p(ct))
        return ct

    def trivial(self, pt, big=0):
        ct = np.zeros(self.dim(big) + 1, dtype=self.dtype)
        ct[-1] = int(pt) & ((1 << self.bits) - 1)
        return ct

    def phase(self, ct, big=0):
        ct = np.ascontiguousarray(ct, dtype=self.dtype)
        return int(self.L.orc_phase(self.h, big, _p(ct)))

    def encrypt_batch(self, pts, seed0, big=0):
        pts = np.ascontiguousarray(pts, dtype=np.uint64)
        out = np.zeros((len(pts), self.dim(big) + 1), dtype=self.dtype)
        self.L.orc_encrypt_batch(self.h, big, _p(pts), seed0, len(pts), _p(out))
        return out

    def phase_batch(self, cts, big=0):
        cts = np.ascontiguousarray(cts, dtype=self.dtype).reshape(-1, self.dim(big) + 1)
        out = np.zeros(cts.shape[0], dtype=np.uint64)
        self.L.orc_phase_batch(self.h, big, _p(cts), cts.shape[0], _p(out))
        return out

    def enc_bool_batch(self, bits, seed0):
        e = 1 << (self.bits - 3)
        return self.encrypt_batch(np.where(np.asarray(bits).astype(bool), e, 7 * e), seed0, big=0)

    def dec_bool_batch(self, cts, big=0):
        return self.phase_batch(cts, big) < (1 << (self.bits - 1))

    # ---- boolean encoding (reference src/circuit.rs:27-33, :948)
    def enc_bool(self, v, seed):
        return self.encrypt((1 if v else 7) << (self.bits - 3), seed, big=0)

    def dec_bool(self, ct, big=0):
        return self.phase(ct, big) < (1 << (self.bits - 1))

    # ---- shortint encoding (App. A.8): ciphertexts live under the big key
    def delta(self):
        mc = self.P.msg_mod * self.P.carry_mod
        return 1 << (self.bits - 1 - (mc.bit_length() - 1))

    def enc_shortint(self, m, seed):
        return self.encrypt((m % self.P.msg_mod) * self.delta(), seed, big=1)

    def dec_shortint(self, ct):
        d = self.delta()
        v = ((self.phase(ct, 1) + d // 2) // d) % (2 * self.P.msg_mod * self.P.carry_mod)
        return v % self.P.msg_mod

    def dec_shortint_full(self, ct):
        """message and carry bits (mod msg_mod*carry_mod)"""
        d = self.delta()
        return ((self.phase(ct, 1) + d // 2) // d) % (self.P.msg_mod * self.P.carry_mod)

    def keyswitch(self, ct_big):
        ct_big = np.ascontiguousarray(ct_big, dtype=self.dtype)
        out = np.zeros(self.n + 1, dtype=self.dtype)
        self.L.orc_keyswitch(self.h, _p(ct_big), _p(out))
        return out

    def pbs(self, ct_small, tv, mode=EXACT):
        ct_small = np.ascontiguousarray(ct_small, dtype=self.dtype)
        tv = np.ascontiguousarray(tv, dtype=self.dtype)
        out = np.zeros(self.kN + 1, dtype=self.dtype)
        self.L.orc_pbs(self.h, _p(ct_small), _p(tv), _p(out), mode)
        return out

    def cmux(self, i, a_tilde, acc, mode=EXACT):
        acc = np.array(acc, dtype=self.dtype, copy=True).reshape(-1)
        self.L.orc_cmux(self.h, i, a_tilde, _p(acc), mode)
        return acc

    def modswitch(self, ct_small):
        ct_small = np.ascontiguousarray(ct_small, dtype=self.dtype)
        out = np.zeros(self.n + 1, dtype=np.uint32)
        self.L.orc_modswitch(self.h, _p(ct_small), _p(out))
        return out

    def bool_tv(self):
        tv = np.zeros((self.k + 1, self.N), dtype=self.dtype)
        tv[self.k, :] = 1 << (self.bits - 3)
        return tv.reshape(-1)

    def bool_gate(self, op, a, b=None, c=None, mode=EXACT):
        a = np.ascontiguousarray(a, dtype=self.dtype)
        b = a if b is None else np.ascontiguousarray(b, dtype=self.dtype)
        c = a if c is None else np.ascontiguousarray(c, dtype=self.dtype)
        out = np.zeros(self.n + 1, dtype=self.dtype)
        self.L.orc_bool_gate(self.h, op, _p(a), _p(b), _p(c), _p(out), mode)
        return out

    def bool_linear(self, op, a, b):
        out = np.zeros(self.n + 1, dtype=self.dtype)
        self.L.orc_bool_linear(self.h, op, _p(np.ascontiguousarray(a)), _p(np.ascontiguousarray(b)), _p(out))
        return out

    def shortint_accumulator(self, table):
        t = np.ascontiguousarray(table, dtype=np.uint64)
        tv = np.zeros((self.k + 1) * self.N, dtype=self.dtype)
        self.L.orc_shortint_accumulator(self.h, _p(t), _p(tv))
        return tv

    def shortint_lut(self, ct_big, table, mode=EXACT):
        t = np.ascontiguousarray(table, dtype=np.uint64)
        ct_big = np.ascontiguousarray(ct_big, dtype=self.dtype)
        out = np.zeros(self.kN + 1, dtype=self.dtype)
        self.L.orc_shortint_lut(self.h, _p(ct_big), _p(t), _p(out), mode)
        return out

    def bool_level(self, wires, ops, in0, in1, in2, out, mode=FFT):
        """In-place batched evaluation of one level on the host cores; returns threads used."""
        assert wires.dtype == self.dtype and wires.flags.c_contiguous
        ops = np.ascontiguousarray(ops, dtype=np.uint8)
        a = [np.ascontiguousarray(x, dtype=np.uint32) for x in (in0, in1, in2, out)]
        return self.L.orc_bool_level(self.h, _p(wires), _p(ops), _p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]),
                                     len(ops), mode)


def decompose(bits, base_log, level, x):
    out = (C.c_int64 * level)()
    lib().orc_decompose(bits, base_log, level, int(x), out)
    return list(out)
