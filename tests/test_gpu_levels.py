"""GPU parity of hb_eval_level (the batched replacement of the reference's per-level rayon loop,
src/circuit.rs:524-543 -> Gate::evaluate_encrypted src/gates.rs:241-280) against the CPU oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _plain(E, op, x, y, s):
    return {E.AND: x & y, E.NAND: 1 - (x & y), E.OR: x | y, E.NOR: 1 - (x | y), E.XOR: x ^ y,
            E.XNOR: 1 - (x ^ y), E.MUX: x if s else y, E.NOT: 1 - x, E.COPY: x}[op]


@pytest.mark.parametrize("name,reps", [("toy32", 5), ("boolean_default", 1)])
def test_level_all_gate_types_bit_exact(name, reps, keys_factory, engine_factory, oracle_mod):
    """Truth tables of tests/gates_test.rs:9-108 as one batched level; ciphertexts bit-identical to the
    oracle's exact-integer back-end (u32 torus), decryptions equal to the plaintext gate."""
    from helm_b200 import engine as E
    K, eng = keys_factory(name), engine_factory(name)
    ops = [E.AND, E.NAND, E.OR, E.NOR, E.XOR, E.XNOR, E.MUX, E.NOT, E.COPY] * reps
    ng = len(ops)
    rng = np.random.default_rng(11)
    bits = rng.integers(0, 2, size=(ng, 3))
    n_in = 3 * ng
    w = eng.wires(n_in + ng)
    cts = K.enc_bool_batch(bits.reshape(-1), 5000)
    w.write(np.arange(n_in, dtype=np.uint32), cts)
    ar = np.arange(ng)
    gates = E.make_gates(ops, 3 * ar, 3 * ar + 1, n_in + ar, in2=3 * ar + 2)
    eng.eval_level(w, gates)
    eng.sync()
    got = w.read(n_in + ar.astype(np.uint32))
    for g, op in enumerate(ops):
        want = K.bool_gate(op, cts[3 * g], cts[3 * g + 1], cts[3 * g + 2], mode=oracle_mod.EXACT)
        assert np.array_equal(got[g], want), "gate %d op %d differs" % (g, op)
        x, y, s = (int(v) for v in bits[g])
        assert K.dec_bool(got[g]) == bool(_plain(E, op, x, y, s))
    w.close()


def test_two_levels_chain_and_row_io(keys_factory, engine_factory):
    """level 2 consumes level 1's outputs straight from the device-resident wire array; row-range IO."""
    from helm_b200 import engine as E
    K, eng = keys_factory("toy32"), engine_factory("toy32")
    W = 70                                   # ragged vs the 2-gate PBS tile and the 16-gate KS tile
    rng = np.random.default_rng(3)
    b0 = rng.integers(0, 2, size=W)
    w = eng.wires(3 * W)
    w.write_rows(0, K.enc_bool_batch(b0, 900))
    cur = b0
    for lvl in (1, 2):
        ops = rng.integers(0, 6, size=W).astype(np.uint8)
        i0 = rng.integers(0, W, size=W); i1 = rng.integers(0, W, size=W)
        eng.eval_level(w, E.make_gates(ops, (lvl - 1) * W + i0, (lvl - 1) * W + i1, lvl * W + np.arange(W)))
        cur = np.array([_plain(E, int(o), int(cur[a]), int(cur[b]), 0) for o, a, b in zip(ops, i0, i1)])
    eng.sync()
    out = np.zeros((W, eng.ct_words), dtype=eng.dtype)
    w.read_rows(2 * W, out)
    assert np.array_equal(K.dec_bool_batch(out), cur.astype(bool))
    w.close()


def test_empty_level_and_errors(keys_factory, engine_factory):
    import helm_b200
    from helm_b200 import engine as E
    eng = engine_factory("toy32")
    w = eng.wires(8)
    eng.eval_level(w, E.make_gates([], [], [], []))          # empty level is a no-op
    with pytest.raises(helm_b200.EngineError, match="out of range"):
        eng.eval_level(w, E.make_gates([E.AND], [0], [1], [99]))
    with pytest.raises(helm_b200.EngineError):
        eng.eval_level(w, E.make_gates([E.LUT], [0], [1], [2]))   # LUT needs shortint (order 1) parameters
    w.close()


@pytest.mark.parametrize("name", ["toy64", "toy64b"])
def test_shortint_lut_level(name, keys_factory, engine_factory, oracle_mod):
    """LUT mode (src/gates.rs:754-785, 2-input path): KS -> PBS with the boxed accumulator.  u64 torus through
    the f64 FFT is not bit-exact, so: decryptions equal the table, and the phase agrees with the oracle's
    exact-integer result to within the scheme's own output noise (far less than the encoding step)."""
    from helm_b200 import engine as E
    K, eng = keys_factory(name), engine_factory(name)
    tables = [0x8, 0x6, 0xE, 0x1, 0x9, 0x7]
    ng = 4 * len(tables)
    w = eng.wires(2 * ng + ng)
    bits = np.array([[(g >> 1) & 1, g & 1] for g in range(ng)])
    cts = np.stack([K.enc_shortint(int(b), 700 + i) for i, b in enumerate(bits.reshape(-1))])
    w.write(np.arange(2 * ng, dtype=np.uint32), cts)
    ar = np.arange(ng)
    luts = np.array([tables[g // 4] for g in range(ng)], dtype=np.uint64)
    eng.eval_level(w, E.make_gates([E.LUT] * ng, 2 * ar, 2 * ar + 1, 2 * ng + ar, lut=luts, arity=[2] * ng))
    eng.sync()
    got = w.read((2 * ng + ar).astype(np.uint32))
    mc = K.P.msg_mod * K.P.carry_mod
    for g in range(ng):
        a, b = (int(v) for v in bits[g])
        want_bit = (int(luts[g]) >> (a * 2 + b)) & 1
        assert K.dec_shortint(got[g]) == want_bit, "gate %d" % g
        lin = (cts[2 * g] * K.dtype(2) + cts[2 * g + 1]).astype(K.dtype)
        table = [(int(luts[g]) >> x) & 1 if x < 4 else 0 for x in range(mc)]
        ref = K.shortint_lut(lin, table, mode=oracle_mod.EXACT)
        d = (K.phase(got[g], 1) - K.phase(ref, 1)) % (1 << 64)
        d = min(d, (1 << 64) - d)
        # two valid samples of the PBS output noise (digit roundings differ), far inside the decoding margin
        assert d < K.delta() >> 6, "phase differs from the exact oracle by 2^%.1f" % np.log2(max(d, 1))
    w.close()


def test_wide_level_throughput_kernel_bit_exact(keys_factory, engine_factory, oracle_mod):
    """A level wider than the narrow-batch threshold (3 x SMs) runs the throughput kernel (pbs16_kernel: half-warp per
    polynomial, TMEM-resident twiddles, split-barrier MAC); its ciphertexts must equal the oracle's exact-integer
    back-end word for word, including the ragged last CTA (width not a multiple of 4 gates)."""
    from helm_b200 import engine as E
    K, eng = keys_factory("toy32"), engine_factory("toy32")
    W = 601
    rng = np.random.default_rng(17)
    bits = rng.integers(0, 2, size=(W, 3))
    ops = rng.choice([E.AND, E.NAND, E.OR, E.NOR, E.XOR, E.XNOR, E.MUX], size=W).astype(np.uint8)
    w = eng.wires(4 * W)
    cts = K.enc_bool_batch(bits.reshape(-1), 77000)
    w.write_rows(0, cts)
    ar = np.arange(W)
    before = eng.stats()
    eng.eval_level(w, E.make_gates(ops, 3 * ar, 3 * ar + 1, 3 * W + ar, in2=3 * ar + 2))
    eng.sync()
    assert eng.stats()["pbs_count"] - before["pbs_count"] == W + int((ops == E.MUX).sum())
    got = np.zeros((W, eng.ct_words), dtype=eng.dtype)
    w.read_rows(3 * W, got)
    for g in range(W):
        want = K.bool_gate(int(ops[g]), cts[3 * g], cts[3 * g + 1], cts[3 * g + 2], mode=oracle_mod.EXACT)
        assert np.array_equal(got[g], want), "gate %d op %d differs" % (g, ops[g])
        x, y, s = (int(v) for v in bits[g])
        assert K.dec_bool(got[g]) == bool(_plain(E, int(ops[g]), x, y, s))
    w.close()


def test_async_row_io_pipeline_matches_synchronous(keys_factory, engine_factory):
    """hb_wires_write_rows_async / hb_wires_read_rows_async / hb_io_sync: three pipelined steps on disjoint row blocks give the same
    ciphertexts as the synchronous calls (the copy stream orders itself against the compute stream with events)."""
    import torch
    from helm_b200 import engine as E
    K, eng = keys_factory("toy32"), engine_factory("toy32")
    W = 300
    rng = np.random.default_rng(29)
    w = eng.wires(6 * W)
    tdt = torch.int32
    ins, outs, gates, want = [], [], [], []
    for s in range(3):
        bits = rng.integers(0, 2, size=W)
        h = torch.empty((W, eng.ct_words), dtype=tdt).pin_memory()
        h.numpy().view(eng.dtype)[:] = K.enc_bool_batch(bits, 8800 + 1000 * s)
        ins.append(h)
        outs.append(torch.zeros((W, eng.ct_words), dtype=tdt).pin_memory())
        ops = rng.integers(0, 6, size=W).astype(np.uint8)
        i0, i1 = rng.integers(0, W, size=W), rng.integers(0, W, size=W)
        gates.append(E.make_gates(ops, 2 * s * W + i0, 2 * s * W + i1, (2 * s + 1) * W + np.arange(W)))
        want.append(np.array([_plain(E, int(o), int(bits[a]), int(bits[b]), 0) for o, a, b in zip(ops, i0, i1)], dtype=bool))
    for s in range(3):
        w.write_rows_async(2 * s * W, ins[s].numpy().view(eng.dtype))
        eng.eval_level(w, gates[s])
        w.read_rows_async((2 * s + 1) * W, outs[s].numpy().view(eng.dtype))
    eng.io_sync()
    eng.sync()
    for s in range(3):
        got = outs[s].numpy().view(eng.dtype)
        assert np.array_equal(K.dec_bool_batch(got), want[s]), "step %d" % s
        ref = np.zeros((W, eng.ct_words), dtype=eng.dtype)
        w.read_rows((2 * s + 1) * W, ref)
        assert np.array_equal(got, ref)
    w.close()


@pytest.mark.parametrize("name,width", [("toy32", 48), ("boolean_default", 40)])
def test_cluster_kernel_repeatable_and_bit_exact(name, width, keys_factory, engine_factory, oracle_mod):
    """The narrowest levels run one bootstrap per cluster of K+1 SMs (pbs3c_kernel: partial spectra exchanged through distributed
    shared memory with byte-counted mbarriers and double-buffered receive slots).  compute-sanitizer is not available on the pool, so
    the exchange protocol is checked by determinism: 25 launches of the same level, including MUX gates (two bootstraps each) and a
    ragged width, must give the same words every time, equal to the oracle's exact-integer back-end."""
    from helm_b200 import engine as E
    K, eng = keys_factory(name), engine_factory(name)
    W = width
    rng = np.random.default_rng(31)
    bits = rng.integers(0, 2, size=(W, 3))
    ops = rng.choice([E.AND, E.NAND, E.OR, E.NOR, E.XOR, E.XNOR, E.MUX], size=W).astype(np.uint8)
    ops[: W // 8] = E.MUX
    n_pbs = W + int((ops == E.MUX).sum())
    while n_pbs > 48:                                # keep the launch inside the cluster kernel's batch limit
        ops[np.nonzero(ops == E.MUX)[0][0]] = E.AND
        n_pbs -= 1
    cts = K.enc_bool_batch(bits.reshape(-1), 60600)
    w = eng.wires(4 * W)
    w.write_rows(0, cts)
    ar = np.arange(W)
    gates = E.make_gates(ops, 3 * ar, 3 * ar + 1, 3 * W + ar, in2=3 * ar + 2)
    first = None
    for rep in range(25):
        eng.eval_level(w, gates)
        eng.sync()
        got = np.zeros((W, eng.ct_words), dtype=eng.dtype)
        w.read_rows(3 * W, got)
        if first is None:
            first = got
        assert np.array_equal(got, first), "launch %d differs from launch 0" % rep
    ng = min(W, 12 if name == "boolean_default" else W)
    for g in range(ng):
        want = K.bool_gate(int(ops[g]), cts[3 * g], cts[3 * g + 1], cts[3 * g + 2], mode=oracle_mod.EXACT)
        assert np.array_equal(first[g], want), "gate %d op %d differs from the oracle" % (g, ops[g])
    w.close()


def test_negated_input_flags_equal_explicit_not(keys_factory, engine_factory, oracle_mod):
    """hb_gate.flags bit s = input s taken negated (a NOT folded into its consumer): the ciphertext must equal, word for word, the
    oracle's gate on the explicitly negated input -- every gate type, every flag combination."""
    from helm_b200 import engine as E
    K, eng = keys_factory("toy32"), engine_factory("toy32")
    ops, flags = [], []
    for op in (E.AND, E.NAND, E.OR, E.NOR, E.XOR, E.XNOR):
        for f in range(4):
            ops.append(op); flags.append(f)
    for f in range(8):
        ops.append(E.MUX); flags.append(f)
    ng = len(ops)
    rng = np.random.default_rng(41)
    bits = rng.integers(0, 2, size=(ng, 3))
    cts = K.enc_bool_batch(bits.reshape(-1), 31300)
    w = eng.wires(4 * ng)
    w.write_rows(0, cts)
    ar = np.arange(ng)
    eng.eval_level(w, E.make_gates(ops, 3 * ar, 3 * ar + 1, 3 * ng + ar, in2=3 * ar + 2, flags=flags))
    eng.sync()
    got = np.zeros((ng, eng.ct_words), dtype=eng.dtype)
    w.read_rows(3 * ng, got)
    neg = lambda c: (np.zeros_like(c) - c).astype(K.dtype)
    for g in range(ng):
        a, b, c = (neg(cts[3 * g + s]) if (flags[g] >> s) & 1 else cts[3 * g + s] for s in range(3))
        want = K.bool_gate(int(ops[g]), a, b, c, mode=oracle_mod.EXACT)
        assert np.array_equal(got[g], want), "op %d flags %d" % (ops[g], flags[g])
        x, y, s = (int(bits[g, q]) ^ ((flags[g] >> q) & 1) for q in range(3))
        assert K.dec_bool(got[g]) == bool(_plain(E, int(ops[g]), x, y, s))
    w.close()


def test_cuda_graph_replay_of_a_batch_sequence(keys_factory, engine_factory, oracle_mod):
    """hb_graph_begin / hb_graph_end / hb_graph_launch: two dependent levels (a narrow one on the cluster kernel and a wide
    one on the throughput kernels) recorded once and replayed three times on fresh inputs give, each time, the ciphertexts of the
    eager evaluation -- i.e. of the oracle."""
    from helm_b200 import engine as E
    K, eng = keys_factory("toy32"), engine_factory("toy32")
    W1, W2 = 40, 300
    rng = np.random.default_rng(53)
    w = eng.wires(W1 + W1 + W2)
    ops1 = rng.integers(0, 6, size=W1).astype(np.uint8)
    a1, b1 = rng.integers(0, W1, size=W1), rng.integers(0, W1, size=W1)
    ops2 = rng.integers(0, 6, size=W2).astype(np.uint8)
    a2, b2 = W1 + rng.integers(0, W1, size=W2), W1 + rng.integers(0, W1, size=W2)
    g1 = E.make_gates(ops1, a1, b1, W1 + np.arange(W1))
    g2 = E.make_gates(ops2, a2, b2, 2 * W1 + np.arange(W2))

    def eager(cts):
        w.write_rows(0, cts)
        eng.eval_level(w, g1); eng.eval_level(w, g2); eng.sync()
        out = np.zeros((W2, eng.ct_words), dtype=eng.dtype)
        w.read_rows(2 * W1, out)
        return out
    inputs = [K.enc_bool_batch(rng.integers(0, 2, size=W1), 71000 + 100 * i) for i in range(4)]
    want = [eager(c) for c in inputs]                      # also sizes the scratch buffers
    before = eng.stats()
    eng.graph_begin(1 << 20)
    eng.eval_level(w, g1); eng.eval_level(w, g2); eng.sync()     # recorded, not run
    g = eng.graph_end()
    assert eng.stats()["pbs_count"] == before["pbs_count"]
    for i in (1, 2, 3):
        w.write_rows(0, inputs[i])
        eng.graph_launch(g)
        eng.sync()
        out = np.zeros((W2, eng.ct_words), dtype=eng.dtype)
        w.read_rows(2 * W1, out)
        assert np.array_equal(out, want[i]), "replay %d differs from the eager evaluation" % i
    assert eng.stats()["pbs_count"] - before["pbs_count"] == 3 * (W1 + W2)
    eng.graph_destroy(g)
    cpu = np.concatenate([inputs[1], np.zeros((W1 + W2, eng.ct_words), dtype=eng.dtype)])
    K.bool_level(cpu, ops1, a1, b1, a1, W1 + np.arange(W1), mode=oracle_mod.EXACT)
    K.bool_level(cpu, ops2[:16], a2[:16], b2[:16], a2[:16], 2 * W1 + np.arange(16), mode=oracle_mod.EXACT)
    assert np.array_equal(want[1][:16], cpu[2 * W1:2 * W1 + 16])
    w.close()


def test_io_stream_and_fence_order_caller_work_before_evaluation(keys_factory, engine_factory):
    """hb_io_stream / hb_io_fence: work the caller enqueues on the upload stream after an asynchronous upload (in bench.py: the
    all-gather of the ranks' input shares; here: a device copy of the uploaded rows into another block of an attached wire array)
    happens before every evaluation issued after the fence."""
    import torch
    from helm_b200 import engine as E
    K, eng = keys_factory("toy32"), engine_factory("toy32")
    W = 200
    dev = torch.device("cuda", 0)
    wt = torch.zeros((3 * W, eng.stride), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    w = eng.wires(3 * W, device_ptr=wt.data_ptr())
    io = torch.cuda.ExternalStream(eng.io_stream(), device=dev)
    rng = np.random.default_rng(61)
    for rep in range(3):
        bits = rng.integers(0, 2, size=W)
        h = torch.empty((W, eng.ct_words), dtype=torch.int32).pin_memory()
        h.numpy().view(eng.dtype)[:] = K.enc_bool_batch(bits, 9100 + rep)
        w.write_rows_async(0, h.numpy().view(eng.dtype))                   # block 0 <- host
        with torch.cuda.stream(io):
            wt[W:2 * W].copy_(wt[0:W], non_blocking=True)                  # caller work on the upload stream: block 1 <- block 0
        eng.io_fence()
        ops = rng.integers(0, 6, size=W).astype(np.uint8)
        i0, i1 = rng.integers(0, W, size=W), rng.integers(0, W, size=W)
        eng.eval_level(w, E.make_gates(ops, W + i0, W + i1, 2 * W + np.arange(W)))   # reads block 1
        eng.sync()
        eng.io_sync()
        out = np.zeros((W, eng.ct_words), dtype=eng.dtype)
        w.read_rows(2 * W, out)
        want = np.array([_plain(E, int(o), int(bits[a]), int(bits[b]), 0) for o, a, b in zip(ops, i0, i1)], dtype=bool)
        assert np.array_equal(K.dec_bool_batch(out), want), "repetition %d" % rep
    w.close()
