#!/usr/bin/env python
"""bench.py -- PBS/s of batched gate bootstraps on the synthetic levelized netlist (BASELINE.json configs[4]:
65 536 two-input gates per level, depth 32, tfhe boolean DEFAULT_PARAMETERS), SURVEY.md section 8(d) item 5.

A "step" is one netlist level: ONE hb_eval_level call = 65 536 gate bootstraps (fused linear part -> blind
rotation -> sample extract -> keyswitch), i.e. one pass of the hot path over one batch.  Step s evaluates level
(s mod 32)+1, reading the wires of level (s mod 32) and writing the next block of the device-resident wire array.
With N > 1 GPUs every level is sharded across the ranks (W/N gates each; BSK/KSK replicated) and the level's
output ciphertexts are exchanged with one NCCL all-gather before the next level ("strong" scaling: the total
work per step is fixed).

  value    device-resident throughput (inputs already in HBM), CUDA events on the engine's stream, max over ranks
  e2e      the same levels through the C ABI with HOST buffers: hb_wires_write_rows (pinned host -> device) of the
           level's input wires, hb_eval_level, hb_wires_read_rows of the level's outputs, all inside the timed region
  roofline the PBS kernel: algorithmic FP64 flops per launch (141.95 MFLOP per PBS) / its CUDA-event
           duration, against the DFMA peak measured in this run by a register-resident FMA loop
  cpu_baseline / --impl reference: the CPU restatement of the reference's tfhe-rs path (oracle/, f64-FFT mode, one
           gate per thread like rayon par_iter) on the host cores, on a bounded sample of the same level.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W_DEFAULT, DEPTH = 65536, 32
PARAM_SET = "boolean_default"
F_CMUX = 196608.0            # flops per CMUX step, boolean default


def ncu_traffic(kernel_prefix, width):
    """roofline.traffic: dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the dominant kernel at this width, from
    the newest committed record profiles/*_traffic.json (written by scripts/ncu_traffic.py out of an `ncu --set full` capture).
    None when no record matches the kernel and width that were just measured -- a stale constant is worse than none."""
    import glob
    best = None
    for f in sorted(glob.glob(os.path.join(ROOT, "profiles", "*_traffic.json"))):
        try:
            r = json.load(open(f))
        except Exception:
            continue
        if str(r.get("kernel", "")).startswith(kernel_prefix) and int(r.get("gates_per_launch", -1)) == int(width):
            best = (r, os.path.relpath(f, ROOT))
    if best is None:
        return None, None
    r, f = best
    return float(r["dram_bytes_read"]) + float(r["dram_bytes_write"]), "%s (%s)" % (f, r.get("source", "ncu --set full"))
GATE_OPS = 6                 # AND NAND OR NOR XOR XNOR


# ---------------------------------------------------------------- deterministic workload (splitmix64)
def splitmix64(seed, count):
    x = (np.uint64(seed) + np.arange(1, count + 1, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15))
    z = x
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def synth_level(level, W):
    """gate (level, j): two wires drawn uniformly from level-1, type uniform over the 6 two-input gates (seed 1)."""
    r = splitmix64(1 + 1000003 * level, 3 * W)
    in0 = (r[0::3] % np.uint64(W)).astype(np.uint32)
    in1 = (r[1::3] % np.uint64(W)).astype(np.uint32)
    ops = (r[2::3] % np.uint64(GATE_OPS)).astype(np.uint8)
    return ops, in0, in1


def plain_gate(ops, a, b):
    a = a.astype(bool); b = b.astype(bool)
    res = np.zeros(len(ops), dtype=bool)
    for op, f in enumerate((a & b, ~(a & b), a | b, ~(a | b), a ^ b, ~(a ^ b))):
        res = np.where(ops == op, f, res)
    return res


# ---------------------------------------------------------------- clocks sampler
class Clocks:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(device), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "200"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 6:
                continue
            try:
                sm.append(float(c[0])); mx.append(float(c[1]))
            except ValueError:
                continue
            for nm, v in zip(names, c[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                   "samples": len(sm)}
        return out


# ---------------------------------------------------------------- CPU arm (oracle = restated reference path)
def cpu_level_sample(K, O, wires0, level_gates, count, mode):
    ops, in0, in1 = level_gates
    nw = wires0.shape[0]
    buf = np.concatenate([wires0, np.zeros((count, wires0.shape[1]), dtype=wires0.dtype)])
    t = time.perf_counter()
    threads = K.bool_level(buf, ops[:count], in0[:count], in1[:count], in0[:count], nw + np.arange(count), mode=mode)
    return time.perf_counter() - t, threads, buf[nw:]


def run_reference(args):
    """--impl reference: the reference's CPU path (restated; tfhe-rs itself cannot be built here) on the host cores.
    The loop it stands for is gates.par_iter_mut() over one level, /root/reference/src/circuit.rs:531.
    Under torch.distributed.run (N > 1) only rank 0 works, with every online core: the launcher exports OMP_NUM_THREADS=1
    to its workers, so the OpenMP team is set explicitly, and the per-step sample is sized from a calibration step so that
    the whole arm takes about REF_BUDGET_S seconds whatever the core count."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle as O
    threads = O.set_threads(0)
    K = O.Keys(PARAM_SET, seed=3)
    W = args.width
    bits0 = (splitmix64(2, W) & np.uint64(1)).astype(np.uint8)

    def step(s, sample):
        lvl = s % DEPTH + 1
        ops, in0, in1 = synth_level(lvl, W)
        need = np.unique(np.concatenate([in0[:sample], in1[:sample]]))          # only the wires the sample reads
        remap = np.zeros(W, dtype=np.uint32); remap[need] = np.arange(len(need), dtype=np.uint32)
        w0 = K.enc_bool_batch(bits0[need], 10_000 + 7 * s)
        dt, used, out = cpu_level_sample(K, O, w0, (ops, remap[in0], remap[in1]), sample, O.FFT)
        want = plain_gate(ops[:sample], bits0[in0[:sample]], bits0[in1[:sample]])
        assert np.array_equal(K.dec_bool_batch(out), want), "reference arm: wrong gate outputs"
        return dt, used

    REF_BUDGET_S = 45.0
    dt, used = step(0, 2 * threads)                       # calibration (also builds the Fourier key): untimed
    dt, used = step(0, 4 * threads)
    rate = 4 * threads / dt
    if args.cpu_gates:
        sample = min(W, args.cpu_gates)
    else:
        sample = int(rate * REF_BUDGET_S / max(1, args.warmup + args.steps))
        sample = max(threads, min(W, (sample // threads) * threads))
    times = []
    for s in range(args.warmup + args.steps):
        dt, used = step(s, sample)
        if s >= args.warmup:
            times.append(dt)
    tot = float(sum(times))
    v = sample * len(times) / tot
    line = {"impl": "reference", "metric": "PBS/s (batched gate bootstraps)", "value": v, "unit": "PBS/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot / len(times),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, 1),
            "cpu_baseline": {"value": v, "unit": "PBS/s", "cores": int(used), "kind": "port",
                             "sample": "%d gates of the level per step (of %d), oracle f64-FFT, OpenMP one gate per thread "
                                       "(team of %d set explicitly; host has %d cores); tfhe-rs itself is unavailable on this box -- "
                                       "its hand-vectorised FFT is typically 2-3x faster per core than this port"
                                       % (sample, W, used, os.cpu_count() or 1)},
            "e2e": {"value": v, "unit": "PBS/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit_line(line)


_REAL_STDOUT = None


def emit_line(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_REAL_STDOUT, data)


def workload_config(args, n):
    if getattr(args, "single_process", False) and args.gpus > 1:
        n = args.gpus
    return {"workload": "synthetic levelized netlist, %d two-input gates/level, depth %d (BASELINE.json configs[4])"
                        % (args.width, DEPTH),
            "params": "tfhe boolean DEFAULT_PARAMETERS n=722 k=2 N=512 l=3 B=2^6 ks l=4 B=2^3 (u32 torus)",
            "step": "one level = one hb_eval_level batch of %d gate bootstraps (PBS + keyswitch each)" % args.width,
            "sharding": ("level sharded over %d GPU(s), keys replicated, %s" % (
                            n, "one process: hb_ctx_create_multi, outputs stored into peer replicas over NVLink + per-level events"
                            if getattr(args, "single_process", False) else "one rank per GPU, NCCL all-gather of outputs per level"))
                        if n > 1 else "single GPU",
            "l2": "inputs larger than L2: each level reads/writes a different 190 MB block of the 6.3 GB wire array"}


def hbm_peak_gbs():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0        # fallback stated in B200_PROFILING.md


def iscas_latency(name, device, params=PARAM_SET):
    """End-to-end latency of one ISCAS'85 netlist (all-false inputs, the reference README invocation) in gates mode
    (boolean parameters) or LUT mode (a `_lut` netlist under a shortint parameter set): evaluate = every level as one
    hb_eval_level batch, wires resident; outputs checked against the plaintext netlist."""
    from helm_b200 import host as H
    path = os.path.join(ROOT, "tests", "golden", "yosys", name + ".v")
    if not os.path.exists(path):
        return None
    circ = H.Circuit(text=H.preprocess_yosys(open(path).read(), exact_consts=True))
    sess = H.Session(params, device=device, seed=5)
    want = circ.evaluate()
    best = None
    lv0 = sess.stats()["levels"]
    for rep in range(8):
        got, secs = sess.run(circ)
        assert all(bool(got[k]) == bool(want[k]) for k in got), "c7552: decrypted outputs differ from the plaintext netlist"
        best = secs if best is None or secs[1] < best[1] else best
    widths = circ.level_widths()
    pbs = circ.count(H.Circuit.PBS_GATES) + sum(1 for g in circ.gates() if g[1] == "mux")
    batches = (sess.stats()["levels"] - lv0) // 8          # hb_eval_level calls per evaluation: NOT / BUF folded into their consumers
    sess.close()
    return {"netlist": name, "params": params, "levels": len(widths), "batches": int(batches), "max_width": max(widths), "pbs": pbs,
            "evaluate_ms": 1e3 * best[1], "encrypt_ms": 1e3 * best[0], "decrypt_ms": 1e3 * best[2],
            "pbs_per_s": pbs / best[1], "inputs": "all false", "outputs_verified": True}


def iscas_cpu_latency(name):
    """The same netlist, all-false inputs, evaluated level by level by the CPU oracle (f64-FFT mode, OpenMP one gate per
    thread = the reference's rayon loop over a level, src/circuit.rs:524-543): the CPU time beside `evaluate_ms`."""
    from helm_b200 import host as H
    from oracle import oracle as O
    path = os.path.join(ROOT, "tests", "golden", "yosys", name + ".v")
    circ = H.Circuit(text=H.preprocess_yosys(open(path).read(), exact_consts=True))
    K = O.Keys(PARAM_SET, seed=5)
    gates = circ.gates()
    want = circ.evaluate()
    names = sorted({g[3] for g in gates} | {w for g in gates for w in g[4]})
    idx = {w: i for i, w in enumerate(names)}
    wires = np.zeros((len(names), K.params["n"] + 1), dtype=K.dtype)
    driven = {g[3] for g in gates}
    prim = [w for w in names if w not in driven]
    wires[[idx[w] for w in prim]] = K.enc_bool_batch(np.zeros(len(prim), dtype=np.uint8), 77)
    code = {"and": O.AND, "nand": O.NAND, "or": O.OR, "nor": O.NOR, "xor": O.XOR, "xnor": O.XNOR, "mux": O.MUX,
            "not": O.NOT, "buf": O.COPY, "dff": O.COPY}
    levels = {}
    for lv, ty, _, out, ins in gates:
        if ty in code:
            levels.setdefault(lv, []).append((code[ty], idx[ins[0]], idx[ins[1]] if len(ins) > 1 else 0,
                                              idx[ins[2]] if len(ins) > 2 else 0, idx[out]))
    K.bool_level(wires, [O.COPY], [0], [0], [0], [0], mode=O.FFT)            # builds the Fourier key outside the timed region
    t = time.perf_counter()
    threads = 1
    for lv in sorted(levels):
        a = np.array(levels[lv], dtype=np.int64)
        threads = K.bool_level(wires, a[:, 0], a[:, 1], a[:, 2], a[:, 3], a[:, 4], mode=O.FFT)
    dt = time.perf_counter() - t
    outs = [w for w in want if w in idx and w in driven]
    dec = K.dec_bool_batch(wires[[idx[w] for w in outs]])
    ok = all(bool(d) == bool(want[w]) for d, w in zip(dec, outs))
    return {"cpu_evaluate_ms": 1e3 * dt, "cpu_threads": int(threads), "cpu_kind": "port (oracle f64-FFT, one gate per thread per level)",
            "cpu_outputs_verified": bool(ok)}


def f_cmux(k, l, N):
    """algorithmic flops of one CMUX step: ((k+1)l + (k+1)) transforms of M = N/2 points + (k+1)^2 l M complex MACs"""
    M = N // 2
    return ((k + 1) * l + (k + 1)) * (5 * M * np.log2(M) + 6 * M) + (k + 1) ** 2 * l * M * 8


def shortint_leg(name, device, fp64_peak, W=16384, steps=3, warmup=2):
    """LUT-mode throughput (the path of lut(), /root/reference/src/gates.rs:754-785, keygen /root/reference/src/bin/helm.rs:301):
    W two-input LUT gates per level on a u64 parameter set, keyswitch -> PBS each, wires resident; PBS kernel time from the
    engine's event ring; fraction of the FP64 peak with the set's own F_pbs (M1C1 86.9 MFLOP, M2C2 194.5 MFLOP)."""
    import helm_b200
    from helm_b200 import engine as E
    from oracle import oracle as O
    K = O.Keys(name, seed=3)
    eng = helm_b200.Engine(K.params, device=device)
    eng.load_keys(K.bsk(), K.ksk())
    rng = np.random.default_rng(12)
    bits0 = rng.integers(0, 2, size=W).astype(np.uint64)
    w = eng.wires(3 * W)
    w.write_rows(0, K.encrypt_batch(bits0 * np.uint64(K.delta()), 660000, big=1))
    tabs = rng.integers(1, 15, size=(2, W)).astype(np.uint64)
    ins = rng.integers(0, W, size=(2, 2, W))
    lv = [E.make_gates([E.LUT] * W, b * W + ins[b, 0], b * W + ins[b, 1], (b + 1) * W + np.arange(W), lut=tabs[b], arity=[2] * W)
          for b in range(2)]
    # block 0 -> 1 -> 2, then 1 -> 2 again and again (level 2's outputs are never fed back, the inputs stay valid bits)
    def step(i):
        eng.eval_level(w, lv[0] if i == 0 else lv[1])
    for i in range(warmup):
        step(i)
    eng.sync()
    st0 = eng.stats()
    t0 = time.perf_counter()
    for i in range(steps):
        step(1)
    eng.sync()
    wall = time.perf_counter() - t0
    st1 = eng.stats()
    pbs_ms = (st1["pbs_ms_total"] - st0["pbs_ms_total"]) / steps
    ks_ms = (st1["ks_ms_total"] - st0["ks_ms_total"]) / steps
    b1 = (tabs[0] >> (2 * bits0[ins[0, 0]] + bits0[ins[0, 1]])) & np.uint64(1)
    b2 = (tabs[1] >> (2 * b1[ins[1, 0]] + b1[ins[1, 1]])) & np.uint64(1)
    out = np.zeros((1024, eng.ct_words), dtype=eng.dtype)
    w.read_rows(2 * W, out)
    d = K.delta()
    dec = ((K.phase_batch(out, big=1) + np.uint64(d // 2)) // np.uint64(d)) % np.uint64(K.P.msg_mod)
    ok = bool(np.array_equal(dec, b2[:1024]))
    if not ok:
        raise SystemExit("bench.py: shortint leg %s decrypts wrong" % name)
    P = K.params
    flops = W * P["n"] * f_cmux(P["k"], P["pbs_level"], P["N"])
    ach = flops / (pbs_ms * 1e-3) / 1e12
    w.close(); eng.close()
    return {"params": name, "luts_per_level": W, "pbs_per_s": W / ((pbs_ms + ks_ms) * 1e-3), "wall_pbs_per_s": W * steps / wall,
            "pbs_kernel_ms": pbs_ms, "ks_kernel_ms": ks_ms, "mflop_per_pbs": flops / W / 1e6, "achieved_tflops": ach,
            "frac_of_fp64_peak": ach / fp64_peak, "outputs_verified": ok}


# ---------------------------------------------------------------- GPU arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--width", type=int, default=W_DEFAULT, help="gates per level (default: the BASELINE config, 65536)")
    ap.add_argument("--cpu-gates", type=int, default=0, help="CPU sample size (gates); default 384 x host threads (cpu_baseline), 64 x per step (--impl reference)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-latency", action="store_true", help="skip the c7552 gates-mode latency leg")
    ap.add_argument("--only-shortint", action="store_true", help="run only the LUT-mode throughput legs (kernel A/B runs)")
    ap.add_argument("--no-shortint", action="store_true", help="skip the LUT-mode (u64 parameter sets) throughput legs")
    ap.add_argument("--single-process", action="store_true",
                    help="drive all --gpus devices from THIS process through hb_ctx_create_multi (level sharding and the NVLink "
                         "exchange inside the engine) instead of one torchrun rank per GPU with NCCL")
    args = ap.parse_args()
    # the contract is ONE JSON line on stdout: libraries (NCCL prints its version banner there) get stderr instead
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import helm_b200
    from helm_b200 import engine as E
    from oracle import oracle as O       # checker + cpu_baseline only; never on the measured GPU path

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    group = list(range(args.gpus)) if (args.single_process and args.gpus > 1) else None
    if group:
        assert world == 1, "--single-process is not launched under torchrun"
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- helm_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from helm_b200 import shard
    W = args.width
    assert W % world == 0, "the bench level width must divide by the number of GPUs"
    g0, Wr = shard.level_slice(W, world, rank)

    if args.only_shortint:
        e0 = helm_b200.Engine(O.PARAM_SETS["toy32"], device=local)
        peak = e0.fp64_peak_tflops()
        e0.close()
        emit_line({"lut_mode_throughput": [shortint_leg(nm, local, peak) for nm in ("shortint_m1c1", "shortint_m2c2")], "fp64_peak": peak})
        return
    K = O.Keys(PARAM_SET, seed=3)
    eng = helm_b200.Engine(K.params, device=local, devices=group)
    eng.load_keys(K.bsk(), K.ksk())
    stride, words = eng.stride, eng.ct_words
    # device wire array: (DEPTH+1) level blocks of W rows; attached so torch/NCCL can address it (one engine-owned replica per
    # device in --single-process mode)
    if group:
        wt = None
        wires = eng.wires((DEPTH + 1) * W)
    else:
        wt = torch.zeros(((DEPTH + 1) * W, stride), dtype=torch.int32, device="cuda")
        wires = eng.wires((DEPTH + 1) * W, device_ptr=wt.data_ptr())
    ext = torch.cuda.ExternalStream(eng.stream(), device=torch.device("cuda", local))

    bits0 = (splitmix64(2, W) & np.uint64(1)).astype(np.uint8)
    ct0 = K.enc_bool_batch(bits0, 1_000_000)                     # level-0 inputs, client side
    wires.write_rows(0, ct0)
    levels = []
    for lvl in range(1, DEPTH + 1):
        ops, in0, in1 = synth_level(lvl, W)
        sl = slice(g0, g0 + Wr)
        levels.append(E.make_gates(ops[sl], (lvl - 1) * W + in0[sl], (lvl - 1) * W + in1[sl],
                                   lvl * W + g0 + np.arange(Wr, dtype=np.uint32)))

    def exchange(lvl):
        if world > 1:
            with torch.cuda.stream(ext):
                shard.exchange_level(dist, wt[lvl * W:(lvl + 1) * W], world, rank)

    def step_resident(s):
        lvl = s % DEPTH + 1
        eng.eval_level(wires, levels[lvl - 1])
        exchange(lvl)

    def barrier():
        eng.sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    fp64_peak = eng.fp64_peak_tflops()

    # ---- device-resident timing.  The clock sampler starts before the warm-up: nvidia-smi's start-up (NVML attach) stalls
    # driver calls for tens of ms, which landed inside the timed region when it was started after the barrier
    # (ms_per_step 576 .. 599 for the same 575 ms of kernels); its samples cover warm-up + timed steps, the same load
    clk = Clocks(local) if rank == 0 else None
    for s in range(args.warmup):
        step_resident(s)
    barrier()
    st0 = eng.stats()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    enqueue_ms = []                                  # host time of each step's enqueue (job build + upload + launches)
    for s in range(args.warmup, args.warmup + args.steps):
        th = time.perf_counter()
        step_resident(s)
        enqueue_ms.append(round(1e3 * (time.perf_counter() - th), 3))
    e1.record(ext)
    barrier()
    clocks = clk.stop() if clk else None
    st1 = eng.stats()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    launches = int(st1["kernel_launches"] - st0["kernel_launches"])
    pbs_ms = (st1["pbs_ms_total"] - st0["pbs_ms_total"]) / args.steps
    ks_ms = (st1["ks_ms_total"] - st0["ks_ms_total"]) / args.steps
    gap_ms = (st1["gap_ms_total"] - st0["gap_ms_total"]) / args.steps

    # ---- correctness of what was just timed: replay the netlist in plaintext and decrypt a sample of the last level
    nsteps = args.warmup + args.steps
    last = (nsteps - 1) % DEPTH + 1
    verified = None
    if nsteps <= DEPTH or True:
        cur = bits0.astype(bool)
        plain = {0: cur}
        for lvl in range(1, min(nsteps, DEPTH) + 1):
            ops, in0, in1 = synth_level(lvl, W)
            cur = plain_gate(ops, cur[in0], cur[in1])
            plain[lvl] = cur
        if last in plain:
            ns = min(Wr, 2048)
            got = np.zeros((ns, words), dtype=eng.dtype)
            wires.read_rows(last * W + g0, got)
            dec = K.dec_bool_batch(got)
            verified = bool(np.array_equal(dec, plain[last][g0:g0 + ns]))
            if not verified:
                raise SystemExit("bench.py: decrypted outputs of level %d differ from the plaintext netlist" % last)

    # ---- end to end through the C ABI with host buffers (pinned), copies inside the timed region
    e2e = None
    if not args.no_e2e:
        # every rank holds 1/N of the level's input ciphertexts in pinned host memory (the rows of its own gate slice of the
        # previous level) and wants its own slice of the outputs back: H2D of W/N rows, all-gather of the inputs over NVLink
        # (each gate reads any wire of the previous level), evaluation, all-gather of the outputs (what the next level needs
        # on every rank), D2H of W/N rows
        hin = torch.empty((Wr, words), dtype=torch.int32).pin_memory()
        hout = torch.empty((Wr, words), dtype=torch.int32).pin_memory()
        hin_np = hin.numpy().view(eng.dtype); hout_np = hout.numpy().view(eng.dtype)
        hin_np[:] = ct0[g0:g0 + Wr]

        hout2 = [hout_np, np.empty_like(hout_np)] if group else None
        if not group:
            hb2 = torch.empty((Wr, words), dtype=torch.int32).pin_memory()
            hout2 = [hout_np, hb2.numpy().view(eng.dtype)]

        io_ext = torch.cuda.ExternalStream(eng.io_stream(), device=torch.device("cuda", local)) if (world > 1 and not group) else None
        # its own communicator: NCCL serialises the collectives of one communicator, and the all-gather of step s+1's inputs must not
        # queue behind the all-gather of step s's outputs (which waits for the evaluation of step s)
        in_group = None
        if io_ext is not None:
            opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)   # its kernels must get SM slots while a PBS kernel runs
            in_group = dist.new_group(pg_options=opts)

        def e2e_level(s):
            # inputs in an even block, outputs in the odd block after it: consecutive steps touch disjoint rows, so the upload of step
            # s+1 and the download of step s-1 (copy stream) may overlap the evaluation of step s (compute stream)
            return 2 * (s % (DEPTH // 2)) + 1

        def step_e2e(s):
            lvl = e2e_level(s)
            if group:                                          # one process, several devices: synchronous row IO
                wires.write_rows((lvl - 1) * W + g0, hin_np)
                eng.eval_level(wires, levels[lvl - 1])
                wires.read_rows(lvl * W + g0, hout2[s & 1])
                return
            wires.write_rows_async((lvl - 1) * W + g0, hin_np)   # H2D: this rank's share of the level's input wires
            if world > 1:                                      # inputs visible on every rank: all-gather on the UPLOAD stream, so that
                with torch.cuda.stream(io_ext):                # it overlaps the previous step's evaluation, then a fence for the compute stream
                    shard.exchange_level(dist, wt[(lvl - 1) * W:lvl * W], world, rank, group=in_group)
                eng.io_fence()
            eng.eval_level(wires, levels[lvl - 1])
            exchange(lvl)                                      # outputs visible on every rank
            wires.read_rows_async(lvl * W + g0, hout2[s & 1])  # D2H: this rank's output ciphertexts
        for s in range(args.warmup):
            step_e2e(s)
        eng.io_sync()
        barrier()
        t0 = time.perf_counter()
        e0.record(ext)
        for s in range(args.steps):
            step_e2e(s)
        eng.io_sync()                                          # the last download has landed in host memory
        e1.record(ext)
        barrier()
        wall = (time.perf_counter() - t0) * 1e3
        t = torch.tensor([max(e0.elapsed_time(e1), wall)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        last_s = args.steps - 1
        hout_np = hout2[last_s & 1]
        ops1, i0, i1 = synth_level(e2e_level(last_s), W)
        ns = min(Wr, 1024)
        ok = np.array_equal(K.dec_bool_batch(hout_np[:ns]), plain_gate(ops1, bits0[i0], bits0[i1])[g0:g0 + ns])
        if not ok:
            raise SystemExit("bench.py: e2e outputs decrypt wrong")
        e2e = {"value": W * args.steps / (float(t.item()) * 1e-3), "unit": "PBS/s",
               "h2d_bytes_per_step": int(world * (hin_np.nbytes + levels[0].nbytes)), "d2h_bytes_per_step": int(world * hout_np.nbytes),
               "per_rank": "each of the %d rank(s) copies 1/%d of the level's inputs up and its own outputs down; byte counts "
                           "are the whole job's" % (world, world),
               "pipeline": "synchronous row IO (one process, several devices)" if group else
                           "hb_wires_write_rows_async / hb_wires_read_rows_async: the copies of neighbouring steps overlap the evaluation "
                           "(copy stream); every step's copies are inside the timed region, which ends after hb_io_sync",
               "ms_per_step": float(t.item()) / args.steps}

    # ---- CPU baseline (rank 0, N = 1 only): bounded sample of level 1 on the host cores
    cpu = None
    if rank == 0 and world == 1 and not group and not args.no_cpu:
        threads = os.cpu_count() or 1
        sample = min(W, args.cpu_gates if args.cpu_gates else 384 * threads)                    # ~15 s of host work
        ops1, i0, i1 = synth_level(1, W)
        need = np.unique(np.concatenate([i0[:sample], i1[:sample]]))
        remap = np.zeros(W, dtype=np.uint32); remap[need] = np.arange(len(need), dtype=np.uint32)
        cpu_level_sample(K, O, ct0[need], (ops1, remap[i0], remap[i1]), min(sample, threads), O.FFT)   # warm (Fourier BSK)
        dt, threads, out = cpu_level_sample(K, O, ct0[need], (ops1, remap[i0], remap[i1]), sample, O.FFT)
        gpu = np.zeros((sample, words), dtype=eng.dtype)
        wires.write_rows(0, ct0)
        eng.eval_level(wires, levels[0]); eng.sync()
        wires.read_rows(W, gpu)
        same_bits = bool(np.array_equal(K.dec_bool_batch(out), K.dec_bool_batch(gpu)))
        cpu = {"value": sample / dt, "unit": "PBS/s", "cores": int(threads), "kind": "port",
               "sample": "%d gates of level 1 (of %d), oracle f64-FFT mode, OpenMP one gate per thread; %.1f s; "
                         "decrypts identical to the GPU outputs: %s" % (sample, W, dt, same_bits)}

    # ---- BASELINE.json metric, second half: c7552 gates-mode end-to-end latency (configs[1]) through the host library
    # (Yosys netlist -> restated preprocessor -> levelizer -> one hb_eval_level per level), N = 1 only
    c7552 = c7552_lut = c880 = None
    if rank == 0 and world == 1 and not group and not args.no_latency:
        c7552 = iscas_latency("c7552", local)
        if c7552 is not None and not args.no_cpu:
            c7552.update(iscas_cpu_latency("c7552"))
        c7552_lut = iscas_latency("c7552_lut_4", local, "shortint_m2c2")
        c880 = iscas_latency("c880", local)
    shortint = None
    if rank == 0 and world == 1 and not group and not args.no_shortint:
        shortint = [shortint_leg(nm, local, fp64_peak) for nm in ("shortint_m1c1", "shortint_m2c2")]

    if rank == 0:
        value = W * args.steps / (ms_total * 1e-3)
        flops_per_launch = (Wr // len(group) if group else Wr) * K.n * F_CMUX      # per device
        ach = flops_per_launch / (pbs_ms * 1e-3) / 1e12 if pbs_ms > 0 else None
        bsk_bytes = K.n * 9 * 3 * 256 * 16
        traffic, traffic_src = ncu_traffic("pbs16_kernel", Wr) if (world == 1 and not group) else (None, None)
        # independent anchor for the in-run DFMA measurement: 64 FP64 lanes per SM x 2 flop x the SM clock sampled under load
        sm_count = torch.cuda.get_device_properties(local).multi_processor_count
        sm_mhz = (clocks or {}).get("sm_mhz") or 0.0
        peak_theory = sm_count * 64 * 2 * sm_mhz * 1e6 / 1e12 if sm_mhz else None
        line = {"metric": "PBS/s (batched gate bootstraps)", "value": value, "unit": "PBS/s", "n_gpus": len(group) if group else world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": workload_config(args, world),
                "roofline": {"bound": "fp64", "kernel": "pbs16_kernel<u32,K=2,L=3,BL=6,G=4> (blind rotation + sample extract)",
                             "achieved": ach, "peak": fp64_peak, "unit": "TFLOP/s",
                             "frac": (ach / fp64_peak) if ach else None,
                             "peak_theoretical": peak_theory,
                             "frac_theoretical": (ach / peak_theory) if (ach and peak_theory) else None,
                             "traffic": traffic, "traffic_source": traffic_src,
                             "peak_source": "measured in this run: register-resident DFMA loop (hb_measure_fp64_peak); "
                                            "MEASURED_PEAKS.json has no FP64 entry.  peak_theoretical = SMs x 64 FP64 lanes x 2 "
                                            "flop x the median SM clock nvidia-smi sampled during the timed region",
                             "flops_per_launch": flops_per_launch, "kernel_ms": pbs_ms, "ks_kernel_ms": ks_ms, "gap_ms": gap_ms, "host_enqueue_ms": enqueue_ms,
                             "hbm_floor_ms": 1e3 * (bsk_bytes + 3.0 * Wr * words * 4) / (hbm_peak_gbs() * 1e9)},
                "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
                "verified": verified, "c7552_gates_mode": c7552, "c880_gates_mode": c880, "c7552_lut_mode": c7552_lut,
                "lut_mode_throughput": shortint}
        if group:
            line["multi"] = {k: st1[k] for k in ("n_devices", "sharded_levels", "push_bytes")}
        emit_line(line)
    wires.close()
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
