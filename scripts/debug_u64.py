import sys, numpy as np
sys.path.insert(0, ".")
from oracle import oracle as O
import helm_b200
for name in ("toy64b", "toy64"):
    base = dict(O.PARAM_SETS[name])
    for n in (1, 2, 4):
        P = dict(base); P["n"] = n
        K = O.Keys(P, seed=3)
        eng = helm_b200.Engine(K.params, device=0); eng.load_keys(K.bsk(), K.ksk())
        mc = K.P.msg_mod * K.P.carry_mod
        tv = K.shortint_accumulator(list(range(mc))).reshape(K.k + 1, K.N)
        for label, ct in (("zero", np.zeros(n + 1, dtype=np.uint64)),
                          ("body", np.array([0] * n + [12345 << 50], dtype=np.uint64)),
                          ("a0=N/2", np.array([1 << 62] + [0] * n, dtype=np.uint64)[:n + 1]),
                          ("rand", np.random.default_rng(n).integers(0, 2 ** 63, size=n + 1, dtype=np.uint64) * np.uint64(2))):
            got = eng.pbs_batch(ct[None, :], tv[K.k][None, :])[0]
            ex = K.pbs(ct, tv.reshape(-1), mode=O.EXACT); ff = K.pbs(ct, tv.reshape(-1), mode=O.FFT)
            d = np.abs((got - ex).astype(np.int64)); d2 = np.abs((ff - ex).astype(np.int64))
            print(name, "n=%d" % n, label, "gpu-vs-exact max 2^%.1f (bad words %d) | cpuFFT-vs-exact 2^%.1f | ms %s"
                  % (np.log2(max(int(d.max()), 1)), int((d > (1 << 44)).sum()), np.log2(max(int(d2.max()), 1)), K.modswitch(ct)))
        eng.close()
