#!/usr/bin/env python
"""Device time of one boolean level (PBS + KS) against its width, for picking the narrow-batch kernel thresholds
(engine.cu launch_pbs).  Run on the GPU box; the strategy is chosen with HB_PBS_SMALL / HB_PBS_SMALL_MAX /
HB_PBS_SMALL2_MAX in the environment.  Lives under tests/ because it takes its keys from the oracle (test infrastructure)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import helm_b200
from helm_b200 import engine as E
import oracle as O

PSET = os.environ.get("SWEEP_PARAMS", "boolean_default")
K = O.Keys(PSET, seed=3)
eng = helm_b200.Engine(K.params, device=0)
eng.load_keys(K.bsk(), K.ksk())
widths = [int(x) for x in sys.argv[1:]] or [64, 148, 180, 296, 444, 592, 740, 1184]
WMAX = max(widths)
rng = np.random.default_rng(1)
bits = rng.integers(0, 2, size=WMAX)
w = eng.wires(2 * WMAX)
if PSET.startswith("shortint"):
    w.write_rows(0, np.stack([K.enc_shortint(int(b), 100 + i) for i, b in enumerate(bits)]))
else:
    w.write_rows(0, K.enc_bool_batch(bits, 100))
print("env", {k: v for k, v in os.environ.items() if k.startswith("HB_")})
for W in widths:
    ops = rng.integers(0, 6, size=W).astype(np.uint8)
    i0 = rng.integers(0, WMAX, size=W); i1 = rng.integers(0, WMAX, size=W)
    g = E.make_gates(ops, i0, i1, WMAX + np.arange(W))
    if PSET.startswith("shortint"):
        g["op"] = E.LUT; g["arity"] = 2; g["lut"] = 0x6
    best = None
    for rep in range(4):
        eng.eval_level(w, g); eng.sync()
        st = eng.stats()
        t = st["pbs_ms"] + st["ks_ms"]
        best = t if best is None or t < best else best
    print("width %5d: %.3f ms (pbs %.3f ks %.3f)" % (W, best, st["pbs_ms"], st["ks_ms"]))
    L = E.lib()
    if hasattr(L, "hb_debug_phase_clocks"):           # -DHB_PHASE_CLOCKS build: critical-path cycles of CTA 0, last launch
        import ctypes as C
        buf = (C.c_double * 16)()
        L.hb_debug_phase_clocks.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        if L.hb_debug_phase_clocks(eng.h, buf) == 0:
            names = ["P1a", "P1b", "P2+bar", "keywait", "P3+bar", "P4", "P5+bar", "looptop", "P1b:load+twist", "P1b:dft", "P1b:tw+store",
                     "P4:load", "P4:dft", "P4:tw+store", "P5:load", "P5:dft"]
            n = K.params["n"]
            if os.environ.get("HB_PBS3", "1") != "0" and W <= 148:
                names = ["P1 decompose", "F1", "F2", "F3+bar", "keywait", "MAC+bar", "I3", "I2", "I1+bar", "top", "-", "-", "-", "-", "-", "-"]
                n = n // max(1, K.params.get("grouping", 0) or 1)
            if os.environ.get("SWEEP_PBS3C"):
                names = ["P1", "F1-F3+bar", "keywait", "MAC+send+bar", "recv wait", "I3-I1+bar", "-", "-", "-", "top", "-", "-", "-", "-", "-", "-"]
            if os.environ.get("SWEEP_PBS16"):
                names = ["A", "B", "wait+store", "C", "load+bar", "D", "lastwait", "O+bar", "E", "F", "top", "rotate-pk", "-", "-", "-", "-"]
            print("   cycles/step: " + "  ".join("%s %.0f" % (nm, v / n) for nm, v in zip(names, buf)) + "  total %.0f" % (sum(buf[:10 if os.environ.get("HB_PBS3", "1") != "0" and W <= 148 else 8]) / n))
