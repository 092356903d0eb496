"""Shared helpers of the GPU circuit tests: run a levelized gate netlist on the engine (one hb_eval_level per level) and on
the CPU oracle with the SAME keys and the SAME input ciphertexts, so results compare word for word."""
import numpy as np


def lower(circ, E):
    """H.Circuit -> (wire index, [hb_gate array per level], primary inputs)."""
    gates = circ.gates()
    names = sorted({g[3] for g in gates} | {w for g in gates for w in g[4]})
    idx = {w: i for i, w in enumerate(names)}
    driven = {g[3] for g in gates}
    prim = [w for w in names if w not in driven]
    levels = {}
    for lv, ty, _, out, ins in gates:
        if ty in ("cone", "czero"):
            continue
        i = [idx[w] for w in ins] + [0] * (3 - len(ins))
        levels.setdefault(lv, []).append((E.OP_NAMES[ty], i[0], i[1], i[2], idx[out]))
    return idx, [np.array(levels[lv], dtype=np.int64) for lv in sorted(levels)], prim


def run_both(circ, K, eng, E, O, inputs, seed, mode):
    """Returns (idx, gpu_wires, cpu_wires): every wire row after one pass, from the engine and from the oracle."""
    idx, levels, prim = lower(circ, E)
    bits = np.array([1 if inputs.get(w, False) else 0 for w in prim], dtype=np.uint8)
    cts = K.enc_bool_batch(bits, seed)
    rows = np.array([idx[w] for w in prim], dtype=np.uint32)
    cpu = np.zeros((len(idx), K.n + 1), dtype=K.dtype)
    cpu[:, -1] = 7 << (K.bits - 3)                      # every wire starts as trivial false (src/circuit.rs:455-458)
    cpu[rows] = cts
    w = eng.wires(len(idx))
    w.set_trivial(np.arange(len(idx), dtype=np.uint32), np.full(len(idx), 7 << (K.bits - 3), dtype=np.uint64))
    w.write(rows, cts)
    for a in levels:
        eng.eval_level(w, E.make_gates(a[:, 0], a[:, 1], a[:, 2], a[:, 4], in2=a[:, 3]))
        K.bool_level(cpu, a[:, 0], a[:, 1], a[:, 2], a[:, 3], a[:, 4], mode=mode)
    eng.sync()
    gpu = w.read(np.arange(len(idx), dtype=np.uint32))
    w.close()
    return idx, gpu, cpu
