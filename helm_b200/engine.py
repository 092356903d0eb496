"""ctypes binding of libhelm_b200.so (include/helm_b200.h).

The library is hand-written CUDA for sm_100a; there is no CPU or PyTorch fallback.  Loading fails
loudly if the shared object has not been built (run `python -c "import __graft_entry__ as g; g.build()"`
or `make -C helm_b200/csrc`), and every compute call fails loudly without a CUDA device.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# HB_LIB: a build variant of the same library (kernel A/B measurements, scripts/ab_variants.sh); never a fallback
LIB_PATH = os.environ.get("HB_LIB") or os.path.join(_HERE, "libhelm_b200.so")

HB_OK = 0
AND, NAND, OR, NOR, XOR, XNOR, MUX, NOT, COPY, LUT = range(10)
OP_NAMES = {"and": AND, "nand": NAND, "or": OR, "nor": NOR, "xor": XOR, "xnor": XNOR, "mux": MUX,
            "not": NOT, "buf": COPY, "dff": COPY, "copy": COPY, "lut": LUT}


class HbParams(C.Structure):
    _fields_ = [(f, C.c_int32) for f in ("n", "k", "N", "pbs_base_log", "pbs_level", "ks_base_log", "ks_level",
                                         "torus_bits", "order", "msg_mod", "carry_mod", "grouping")]


class HbGate(C.Structure):
    _fields_ = [("op", C.c_uint8), ("arity", C.c_uint8), ("flags", C.c_uint16), ("in_", C.c_uint32 * 4),
                ("out", C.c_uint32), ("lut", C.c_uint64)]


class HbStats(C.Structure):
    _fields_ = [("pbs_count", C.c_uint64), ("ks_count", C.c_uint64), ("levels", C.c_uint64),
                ("pbs_ms", C.c_double), ("ks_ms", C.c_double), ("kernel_launches", C.c_uint64),
                ("pbs_ms_total", C.c_double), ("ks_ms_total", C.c_double), ("gap_ms_total", C.c_double),
                ("push_bytes", C.c_uint64), ("n_devices", C.c_uint32), ("sharded_levels", C.c_uint32)]


GATE_DTYPE = np.dtype([("op", np.uint8), ("arity", np.uint8), ("flags", np.uint16), ("in", np.uint32, (4,)),
                       ("out", np.uint32), ("lut", np.uint64)], align=True)
assert GATE_DTYPE.itemsize == C.sizeof(HbGate) == 32
# hb_sop: one shortint block operation (LINEAR = 0 / LUT = 1), include/helm_b200.h
SOP_LINEAR, SOP_LUT = 0, 1
SOP_DTYPE = np.dtype([("kind", np.uint8), ("n_in", np.uint8), ("coef", np.int8, (4,)), ("cst", np.int16),
                      ("in", np.uint32, (4,)), ("out", np.uint32), ("table", np.uint64)], align=True)
assert SOP_DTYPE.itemsize == 40 and SOP_DTYPE.fields["in"][1] == 8 and SOP_DTYPE.fields["table"][1] == 32

EXPORTS = ["hb_ctx_create", "hb_ctx_create_multi", "hb_ctx_destroy", "hb_last_error", "hb_set_params", "hb_load_bsk", "hb_load_ksk",
           "hb_wires_create", "hb_wires_attach", "hb_wires_destroy", "hb_ct_words", "hb_wire_stride",
           "hb_wires_device_ptr", "hb_wires_write", "hb_wires_read", "hb_wires_write_rows", "hb_wires_read_rows", "hb_wires_write_rows_async", "hb_wires_read_rows_async", "hb_io_sync", "hb_io_stream", "hb_io_fence", "hb_wires_set_trivial", "hb_eval_level", "hb_eval_shortint_ops",
           "hb_sync", "hb_stats", "hb_graph_begin", "hb_graph_end", "hb_graph_launch", "hb_graph_destroy", "hb_stream", "hb_pbs_batch", "hb_keyswitch_batch", "hb_keyswitch_batch_noalloc",
           "hb_measure_fp64_peak"]

_LIB = None


class EngineError(RuntimeError):
    pass


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise EngineError("helm_b200: %s is missing -- the CUDA engine is not built and there is no fallback "
                              "(run __graft_entry__.build() or make -C helm_b200/csrc)" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        vp, u32, i32 = C.c_void_p, C.c_uint32, C.c_int32
        L.hb_ctx_create.argtypes = [i32, C.POINTER(vp)]
        L.hb_ctx_create_multi.argtypes = [C.POINTER(i32), i32, u32, C.POINTER(vp)]
        L.hb_ctx_destroy.argtypes = [vp]
        L.hb_last_error.restype = C.c_char_p
        L.hb_last_error.argtypes = [vp]
        L.hb_set_params.argtypes = [vp, C.POINTER(HbParams)]
        L.hb_load_bsk.argtypes = [vp, vp, C.c_size_t]
        L.hb_load_ksk.argtypes = [vp, vp, C.c_size_t]
        L.hb_wires_create.argtypes = [vp, u32, C.POINTER(vp)]
        L.hb_wires_attach.argtypes = [vp, vp, u32, C.POINTER(vp)]
        L.hb_wires_destroy.argtypes = [vp]
        L.hb_ct_words.restype = u32
        L.hb_ct_words.argtypes = [vp]
        L.hb_wire_stride.restype = u32
        L.hb_wire_stride.argtypes = [vp]
        L.hb_wires_device_ptr.restype = vp
        L.hb_wires_device_ptr.argtypes = [vp]
        L.hb_wires_write.argtypes = [vp, vp, u32, vp]
        L.hb_wires_read.argtypes = [vp, vp, u32, vp]
        L.hb_wires_write_rows.argtypes = [vp, u32, u32, vp]
        L.hb_wires_read_rows.argtypes = [vp, u32, u32, vp]
        L.hb_wires_write_rows_async.argtypes = [vp, u32, u32, vp]
        L.hb_wires_read_rows_async.argtypes = [vp, u32, u32, vp]
        L.hb_io_sync.argtypes = [vp]
        L.hb_io_stream.restype = vp
        L.hb_io_stream.argtypes = [vp]
        L.hb_io_fence.argtypes = [vp]
        L.hb_wires_set_trivial.argtypes = [vp, vp, u32, vp]
        L.hb_eval_level.argtypes = [vp, vp, vp, u32]
        L.hb_eval_shortint_ops.argtypes = [vp, vp, vp, u32]
        L.hb_sync.argtypes = [vp]
        L.hb_graph_begin.argtypes = [vp, C.c_size_t]
        L.hb_graph_end.argtypes = [vp, C.POINTER(vp)]
        L.hb_graph_launch.argtypes = [vp]
        L.hb_graph_destroy.argtypes = [vp]
        L.hb_stats.argtypes = [vp, C.POINTER(HbStats)]
        L.hb_stream.restype = vp
        L.hb_stream.argtypes = [vp]
        L.hb_pbs_batch.argtypes = [vp, vp, vp, i32, u32, vp]
        L.hb_keyswitch_batch.argtypes = [vp, vp, u32, vp]
        L.hb_keyswitch_batch_noalloc.argtypes = [vp, vp, u32, vp]
        L.hb_measure_fp64_peak.argtypes = [vp, C.POINTER(C.c_double)]
        _LIB = L
    return _LIB


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def make_gates(ops, in0, in1, out, in2=None, in3=None, lut=None, arity=None, flags=None):
    """Pack parallel arrays into an hb_gate[] numpy array."""
    g = np.zeros(len(ops), dtype=GATE_DTYPE)
    g["op"] = ops
    g["in"][:, 0] = in0
    g["in"][:, 1] = in1
    if in2 is not None:
        g["in"][:, 2] = in2
    if in3 is not None:
        g["in"][:, 3] = in3
    g["out"] = out
    if lut is not None:
        g["lut"] = lut
    if arity is not None:
        g["arity"] = arity
    if flags is not None:
        g["flags"] = flags
    return g


class Engine:
    """One context on one GPU -- or, with devices=[...], on a group of GPUs of this node (hb_ctx_create_multi: keys and wires
    replicated, wide levels sharded, outputs exchanged over NVLink inside the engine): resident Fourier BSK + KSK, batched
    level evaluation."""

    def __init__(self, params, device=0, devices=None, min_gates_per_device=0):
        self.L = lib()
        self.h = C.c_void_p()
        if devices is None:
            st = self.L.hb_ctx_create(device, C.byref(self.h))
        else:
            arr = (C.c_int32 * len(devices))(*devices)
            st = self.L.hb_ctx_create_multi(arr, len(devices), min_gates_per_device, C.byref(self.h))
        if st != HB_OK:
            raise EngineError("hb_ctx_create: %s" % self.L.hb_last_error(None).decode())
        keys = [f for f, _ in HbParams._fields_]
        self.params = {k: int(params.get(k, 0)) for k in keys}
        self.P = HbParams(**self.params)
        self._ck(self.L.hb_set_params(self.h, C.byref(self.P)), "hb_set_params")
        self.dtype = np.uint32 if self.P.torus_bits == 32 else np.uint64
        self.ct_words = self.L.hb_ct_words(self.h)
        self.stride = self.L.hb_wire_stride(self.h)
        self.n, self.kN = self.P.n, self.P.k * self.P.N

    def _ck(self, st, what):
        if st != HB_OK:
            raise EngineError("%s failed (%d): %s" % (what, st, self.L.hb_last_error(self.h).decode()))

    def close(self):
        if self.h:
            self.L.hb_ctx_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def load_keys(self, bsk, ksk):
        bsk = np.ascontiguousarray(bsk, dtype=self.dtype)
        ksk = np.ascontiguousarray(ksk, dtype=self.dtype)
        self._ck(self.L.hb_load_bsk(self.h, _ptr(bsk), bsk.nbytes), "hb_load_bsk")
        self._ck(self.L.hb_load_ksk(self.h, _ptr(ksk), ksk.nbytes), "hb_load_ksk")

    def wires(self, n_wires, device_ptr=None):
        return Wires(self, n_wires, device_ptr)

    def eval_level(self, wires, gates):
        gates = np.ascontiguousarray(gates, dtype=GATE_DTYPE)
        self._ck(self.L.hb_eval_level(self.h, wires.h, _ptr(gates), len(gates)), "hb_eval_level")

    def eval_shortint_ops(self, wires, ops):
        ops = np.ascontiguousarray(ops, dtype=SOP_DTYPE)
        self._ck(self.L.hb_eval_shortint_ops(self.h, wires.h, _ptr(ops), len(ops)), "hb_eval_shortint_ops")

    def sync(self):
        self._ck(self.L.hb_sync(self.h), "hb_sync")

    def graph_begin(self, job_bytes):
        """records (does not run) the eval_level / eval_shortint_ops calls until graph_end(); returns nothing"""
        self._ck(self.L.hb_graph_begin(self.h, job_bytes), "hb_graph_begin")

    def graph_end(self):
        g = C.c_void_p()
        self._ck(self.L.hb_graph_end(self.h, C.byref(g)), "hb_graph_end")
        return g

    def graph_launch(self, g):
        self._ck(self.L.hb_graph_launch(g), "hb_graph_launch")

    def graph_destroy(self, g):
        self.L.hb_graph_destroy(g)

    def io_stream(self):
        return self.L.hb_io_stream(self.h)

    def io_fence(self):
        self._ck(self.L.hb_io_fence(self.h), "hb_io_fence")

    def io_sync(self):
        self._ck(self.L.hb_io_sync(self.h), "hb_io_sync")

    def stats(self):
        s = HbStats()
        self._ck(self.L.hb_stats(self.h, C.byref(s)), "hb_stats")
        return {f: getattr(s, f) for f, _ in HbStats._fields_}

    def stream(self):
        return self.L.hb_stream(self.h)

    def pbs_batch(self, lwe_in, tv_bodies):
        lwe_in = np.ascontiguousarray(lwe_in, dtype=self.dtype).reshape(-1, self.n + 1)
        tv = np.ascontiguousarray(tv_bodies, dtype=self.dtype).reshape(-1, self.P.N)
        shared = 1 if tv.shape[0] == 1 else 0
        assert shared or tv.shape[0] == lwe_in.shape[0]
        out = np.zeros((lwe_in.shape[0], self.kN + 1), dtype=self.dtype)
        self._ck(self.L.hb_pbs_batch(self.h, _ptr(lwe_in), _ptr(tv), shared, lwe_in.shape[0], _ptr(out)), "hb_pbs_batch")
        return out

    def keyswitch_batch(self, lwe_in):
        lwe_in = np.ascontiguousarray(lwe_in, dtype=self.dtype).reshape(-1, self.kN + 1)
        out = np.zeros((lwe_in.shape[0], self.n + 1), dtype=self.dtype)
        self._ck(self.L.hb_keyswitch_batch(self.h, _ptr(lwe_in), lwe_in.shape[0], _ptr(out)), "hb_keyswitch_batch")
        return out

    def keyswitch_batch_noalloc(self, in_dev_ptr, count, out_dev_ptr):
        """hb_keyswitch_batch_noalloc: device pointers in and out, static scratch, no allocation; asynchronous (sync() after)."""
        self._ck(self.L.hb_keyswitch_batch_noalloc(self.h, C.c_void_p(in_dev_ptr), count, C.c_void_p(out_dev_ptr)),
                 "hb_keyswitch_batch_noalloc")

    def fp64_peak_tflops(self):
        v = C.c_double()
        self._ck(self.L.hb_measure_fp64_peak(self.h, C.byref(v)), "hb_measure_fp64_peak")
        return v.value


class Wires:
    """Device-resident wire array (the reference's HashMap<String, Ciphertext>, src/circuit.rs:517-520)."""

    def __init__(self, eng, n_wires, device_ptr=None):
        self.eng, self.n_wires = eng, n_wires
        self.h = C.c_void_p()
        if device_ptr is None:
            eng._ck(eng.L.hb_wires_create(eng.h, n_wires, C.byref(self.h)), "hb_wires_create")
        else:
            eng._ck(eng.L.hb_wires_attach(eng.h, C.c_void_p(device_ptr), n_wires, C.byref(self.h)), "hb_wires_attach")

    def close(self):
        if self.h and self.eng.h:
            self.eng.L.hb_wires_destroy(self.h)
        self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def write(self, idx, cts):
        cts = np.ascontiguousarray(cts, dtype=self.eng.dtype).reshape(-1, self.eng.ct_words)
        if idx is not None:
            idx = np.ascontiguousarray(idx, dtype=np.uint32)
            assert len(idx) == cts.shape[0]
        self.eng._ck(self.eng.L.hb_wires_write(self.h, _ptr(idx), cts.shape[0], _ptr(cts)), "hb_wires_write")

    def read(self, idx=None, count=None, out=None):
        if idx is not None:
            idx = np.ascontiguousarray(idx, dtype=np.uint32)
            count = len(idx)
        elif count is None:
            count = self.n_wires
        if out is None:
            out = np.zeros((count, self.eng.ct_words), dtype=self.eng.dtype)
        self.eng._ck(self.eng.L.hb_wires_read(self.h, _ptr(idx), count, _ptr(out)), "hb_wires_read")
        return out

    def write_rows(self, first, cts):
        """cts: C-contiguous [cnt, ct_words] host array (pinned for full PCIe speed) -> rows [first, first+cnt)."""
        assert cts.dtype == self.eng.dtype and cts.flags.c_contiguous and cts.shape[1] == self.eng.ct_words
        self.eng._ck(self.eng.L.hb_wires_write_rows(self.h, first, cts.shape[0], _ptr(cts)), "hb_wires_write_rows")

    def read_rows(self, first, out):
        assert out.dtype == self.eng.dtype and out.flags.c_contiguous and out.shape[1] == self.eng.ct_words
        self.eng._ck(self.eng.L.hb_wires_read_rows(self.h, first, out.shape[0], _ptr(out)), "hb_wires_read_rows")
        return out

    def write_rows_async(self, first, cts):
        """as write_rows, on the context's copy stream; `cts` (pinned) must stay alive until Engine.io_sync()"""
        assert cts.dtype == self.eng.dtype and cts.flags.c_contiguous and cts.shape[1] == self.eng.ct_words
        self.eng._ck(self.eng.L.hb_wires_write_rows_async(self.h, first, cts.shape[0], _ptr(cts)), "hb_wires_write_rows_async")

    def read_rows_async(self, first, out):
        """as read_rows, on the context's copy stream; `out` is complete after Engine.io_sync()"""
        assert out.dtype == self.eng.dtype and out.flags.c_contiguous and out.shape[1] == self.eng.ct_words
        self.eng._ck(self.eng.L.hb_wires_read_rows_async(self.h, first, out.shape[0], _ptr(out)), "hb_wires_read_rows_async")

    def set_trivial(self, idx, plaintexts):
        idx = np.ascontiguousarray(idx, dtype=np.uint32)
        pt = np.ascontiguousarray(plaintexts, dtype=np.uint64)
        self.eng._ck(self.eng.L.hb_wires_set_trivial(self.h, _ptr(idx), len(idx), _ptr(pt)), "hb_wires_set_trivial")

    def device_ptr(self):
        return self.eng.L.hb_wires_device_ptr(self.h)
