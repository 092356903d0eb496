"""ctypes binding of libhelm_host.so (include/helm_host.h): HELM's circuit surface in C++ over the B200 engine.

Parsing, levelization and plaintext evaluation work anywhere; Session (encrypted evaluation) needs a CUDA device and
fails loudly without one -- there is no CPU fallback.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libhelm_host.so")
EXPORTS = ["hh_last_error", "hh_circuit_load", "hh_circuit_load_text", "hh_circuit_free", "hh_circuit_count",
           "hh_circuit_level_width", "hh_circuit_names", "hh_circuit_describe", "hh_circuit_eval_plain", "hh_read_input_csv", "hh_chacha20_block", "hh_preprocess_yosys", "hh_preprocess_yosys_mode", "hh_preprocess_assign_dict",
           "hh_session_create", "hh_session_create_multi", "hh_session_free", "hh_session_run", "hh_session_run_arith", "hh_session_stats", "hh_session_stats_multi"]
_LIB = None


class HostError(RuntimeError):
    pass


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise HostError("helm_b200: %s is missing (run __graft_entry__.build())" % LIB_PATH)
        C.CDLL(os.path.join(_HERE, "libhelm_b200.so"), mode=C.RTLD_GLOBAL)
        L = C.CDLL(LIB_PATH)
        vp, cp, sz = C.c_void_p, C.c_char_p, C.c_size_t
        L.hh_last_error.restype = cp
        L.hh_circuit_load.argtypes = [cp, C.c_int, C.c_int, C.POINTER(vp)]
        L.hh_circuit_load_text.argtypes = [cp, C.c_int, C.c_int, C.POINTER(vp)]
        L.hh_circuit_free.argtypes = [vp]
        L.hh_circuit_count.restype = C.c_int64
        L.hh_circuit_count.argtypes = [vp, C.c_int]
        L.hh_circuit_level_width.restype = C.c_int64
        L.hh_circuit_level_width.argtypes = [vp, C.c_int]
        L.hh_circuit_describe.argtypes = [vp, vp, sz, C.POINTER(sz)]
        L.hh_circuit_names.argtypes = [vp, C.c_int, vp, sz, C.POINTER(sz)]
        L.hh_circuit_eval_plain.argtypes = [vp, cp, cp, C.c_int, vp, sz, C.POINTER(sz)]
        L.hh_read_input_csv.argtypes = [cp, cp, vp, sz, C.POINTER(sz)]
        L.hh_preprocess_yosys.argtypes = [cp, vp, sz, C.POINTER(sz)]
        L.hh_preprocess_yosys_mode.argtypes = [cp, C.c_int, vp, sz, C.POINTER(sz)]
        L.hh_preprocess_assign_dict.argtypes = [cp, C.c_int, vp, sz, C.POINTER(sz)]
        L.hh_chacha20_block.argtypes = [vp, vp, vp]
        L.hh_chacha20_block.restype = None
        L.hh_session_create.argtypes = [cp, C.c_int, C.c_uint64, vp, vp, vp, sz, vp, sz, C.POINTER(vp)]
        L.hh_session_create_multi.argtypes = [cp, C.POINTER(C.c_int), C.c_int, C.c_uint64, vp, vp, vp, sz, vp, sz, C.POINTER(vp)]
        L.hh_session_stats_multi.argtypes = [vp, C.POINTER(C.c_uint64)]
        L.hh_session_free.argtypes = [vp]
        L.hh_session_run.argtypes = [vp, vp, cp, C.c_int, C.c_int, C.c_int, vp, sz, C.POINTER(sz), C.POINTER(C.c_double)]
        L.hh_session_run_arith.argtypes = [vp, vp, cp, cp, C.c_int, C.c_int, C.c_int, vp, sz, C.POINTER(sz),
                                           C.POINTER(C.c_double), C.POINTER(C.c_uint64)]
        L.hh_session_stats.argtypes = [vp, C.POINTER(C.c_uint64)]
        _LIB = L
    return _LIB


def _ck(rc):
    if rc != 0:
        raise HostError(lib().hh_last_error().decode())


def _text(call, cap=1 << 16):
    """Runs call(buf, cap, needed) growing the buffer until the text fits (the call is repeated then: callers whose call
    does real work pass a cap that holds their largest answer)."""
    while True:
        buf = C.create_string_buffer(cap)
        need = C.c_size_t(0)
        rc = call(buf, cap, C.byref(need))
        if rc != 0 and need.value > cap:      # the text did not fit: grow and call again
            cap = need.value + 16
            continue
        _ck(rc)
        return buf.value.decode()


def _parse(text):
    out = {}
    for line in text.splitlines():
        k, v = line.split("=", 1)
        out[k] = True if v == "true" else False if v == "false" else int(v)
    return out


def _fmt(values):
    if not values:
        return b""
    return "".join("%s=%s\n" % (k, ("true" if v else "false") if isinstance(v, bool) else int(v))
                   for k, v in values.items()).encode()


def preprocess_yosys(text, arith=False, exact_consts=False):
    """Yosys netlist text -> HELM netlist text (reference hdl-benchmarks preprocessor: gate cells, LUTs, multi-bit
    assigns; arith=True for arithmetic designs; exact_consts=True turns the reference's LUT-constant quirk off)."""
    L = lib()
    mode = (1 if arith else 0) | (2 if exact_consts else 0)
    return _text(lambda b, c, n: L.hh_preprocess_yosys_mode(text.encode(), mode, b, c, n))


def preprocess_assign_dict(text, arith=False):
    """Per-bit assign dictionary {dest: source} (reference build_assign_dict)."""
    L = lib()
    txt = _text(lambda b, c, n: L.hh_preprocess_assign_dict(text.encode(), int(arith), b, c, n))
    return dict(line.split("=", 1) for line in txt.splitlines() if line)


def read_input_wires(path, ptxt_type="bool"):
    L = lib()
    return _parse(_text(lambda b, c, n: L.hh_read_input_csv(path.encode(), ptxt_type.encode(), b, c, n)))


class Circuit:
    """read_verilog_file + Circuit::new + sort_circuit + compute_levels (reference src/bin/helm.rs:38-62)."""
    GATES, WIRES, INPUTS, OUTPUTS, DFFS, LEVELS, HAS_LUTS, HAS_ARITH, PBS_GATES, BATCHES = range(10)

    def __init__(self, path=None, text=None, is_arith=False, strict_reference_levels=False):
        self.L = lib()
        self.h = C.c_void_p()
        if path is not None:
            _ck(self.L.hh_circuit_load(path.encode(), int(is_arith), int(strict_reference_levels), C.byref(self.h)))
        else:
            _ck(self.L.hh_circuit_load_text(text.encode(), int(is_arith), int(strict_reference_levels), C.byref(self.h)))

    def __del__(self):
        try:
            if self.h:
                self.L.hh_circuit_free(self.h)
        except Exception:
            pass

    def count(self, what):
        return int(self.L.hh_circuit_count(self.h, what))

    def level_widths(self):
        return [int(self.L.hh_circuit_level_width(self.h, i)) for i in range(self.count(self.LEVELS))]

    def names(self, what):
        """Declared INPUTS (DFF outputs appended, as the reference's parser does), OUTPUTS or DFFS, in file order, unique."""
        txt = _text(lambda b, c, n: self.L.hh_circuit_names(self.h, what, b, c, n))
        return list(dict.fromkeys(x for x in txt.splitlines() if x))

    def inputs(self):
        """Wires read by some gate and driven by none (primary inputs actually used)."""
        g = self.gates()
        driven = {x[3] for x in g}
        return sorted({w for x in g for w in x[4] if w not in driven})

    def gates(self):
        """[(level, type, name, out, [inputs])] in level order."""
        rows = []
        for line in _text(lambda b, c, n: self.L.hh_circuit_describe(self.h, b, c, n)).splitlines():
            lv, ty, name, out, ins = line.split(":", 4)
            rows.append((int(lv), ty, name, out, ins.split(",") if ins else []))
        return rows

    def evaluate(self, inputs=None, ptxt_type="bool", cycles=1):
        """Plaintext evaluation (Circuit::evaluate, src/circuit.rs:348-381): every wire that carries a value."""
        return _parse(_text(lambda b, c, n: self.L.hh_circuit_eval_plain(self.h, _fmt(inputs), ptxt_type.encode(), cycles, b, c, n)))


class Session:
    """Keys + engine context (reference: gen_keys + GateCircuit/LutCircuit::new)."""

    def __init__(self, param_set, device=0, seed=1, raw_keys=None, devices=None):
        """devices=[0, 1, ...]: one session over several GPUs of the node (levels sharded inside the engine)."""
        self.L = lib()
        self.h = C.c_void_p()
        dv = list(devices) if devices is not None else [device]
        arr = (C.c_int * len(dv))(*dv)
        if raw_keys is None:
            _ck(self.L.hh_session_create_multi(param_set.encode(), arr, len(dv), seed, None, None, None, 0, None, 0, C.byref(self.h)))
        else:
            lwe_sk, glwe_sk, bsk, ksk = raw_keys       # numpy arrays (uint8, uint8, torus, torus)
            self._keep = raw_keys
            _ck(self.L.hh_session_create_multi(param_set.encode(), arr, len(dv), seed, lwe_sk.ctypes.data, glwe_sk.ctypes.data,
                                               bsk.ctypes.data, bsk.nbytes, ksk.ctypes.data, ksk.nbytes, C.byref(self.h)))

    def close(self):
        if self.h:
            self.L.hh_session_free(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _open(self):
        if not self.h:
            raise HostError("session is closed")

    def run(self, circuit, inputs=None, cycles=1, all_wires=False, verbose=False):
        """encrypt_inputs -> evaluate_encrypted x cycles -> decrypt_outputs.  Returns (values, seconds[3])."""
        self._open()
        secs = (C.c_double * 3)()
        txt = _text(lambda b, c, n: self.L.hh_session_run(self.h, circuit.h, _fmt(inputs), cycles, int(all_wires),
                                                          int(verbose), b, c, n, secs), cap=1 << 22)
        return _parse(txt), list(secs)

    def run_arith(self, circuit, inputs=None, ptxt_type="u32", cycles=1, all_wires=False, verbose=False):
        """Arithmetic mode (reference ArithCircuit).  Returns (values, seconds[3], {"bootstraps", "batches"})."""
        self._open()
        secs = (C.c_double * 3)()
        cnt = (C.c_uint64 * 2)()
        txt = _text(lambda b, c, n: self.L.hh_session_run_arith(self.h, circuit.h, _fmt(inputs), ptxt_type.encode(), cycles,
                                                                int(all_wires), int(verbose), b, c, n, secs, cnt), cap=1 << 22)
        return _parse(txt), list(secs), {"bootstraps": int(cnt[0]), "batches": int(cnt[1])}

    def stats(self):
        if not self.h:
            raise HostError("session is closed")
        out = (C.c_uint64 * 4)()
        _ck(self.L.hh_session_stats(self.h, out))
        d = dict(zip(("pbs_count", "ks_count", "levels", "kernel_launches"), (int(v) for v in out)))
        m = (C.c_uint64 * 3)()
        _ck(self.L.hh_session_stats_multi(self.h, m))
        d.update(zip(("n_devices", "sharded_levels", "push_bytes"), (int(v) for v in m)))
        return d
