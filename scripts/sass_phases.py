#!/usr/bin/env python
"""Splits an `ncu --page source --print-source sass --csv` dump at BAR.SYNC into phases and prints, per phase,
sampled stall share, executed warp instructions by opcode class and shared-memory wavefronts.
usage: sass_phases.py <sass.csv> [units]   (units = gate-CMUX count to normalise by)"""
import csv, sys, collections, re
rows = list(csv.reader(open(sys.argv[1])))
units = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}
def cls(op):
    o = op.split()[0] if op else ""
    if o.startswith("@"): o = op.split()[1]
    b = o.split(".")[0]
    if b in ("DFMA", "DADD", "DMUL", "DSETP"): return "fp64"
    if b in ("LDS", "STS", "LDSM"): return "smem"
    if b in ("LDG", "STG", "LD", "ST"): return "gmem"
    if b in ("BAR",): return "bar"
    if b in ("IMAD", "IADD3", "LOP3", "SHF", "LEA", "PRMT", "ISETP", "SEL", "IABS", "I2F", "F2I", "VIADD", "IADD", "SGXT", "BMSK", "POPC", "FLO", "I2FP", "F2FP", "IMNMX", "VIMNMX", "LOP", "MOV", "CS2R", "S2R", "SHL", "SHR", "UMOV", "R2UR", "UIADD3", "ULOP3", "USHF", "UIMAD", "ULEA", "R2P", "P2R", "PLOP3", "UISETP", "USEL", "UPLOP3"): return "int/mov:" + b
    return "other:" + b
phases = []; cur = collections.defaultdict(float); cur["_start"] = 0
for n, r in enumerate(rows[2:]):
    if len(r) < len(hdr): continue
    op = r[ix["Source"]].strip(); c = cls(op)
    ie = float(r[ix["Instructions Executed"]] or 0); smp = float(r[ix["# Samples"]] or 0)
    cur["inst"] += ie; cur["samples"] += smp; cur["i:" + c] += ie; cur["wave"] += float(r[ix["L1 Wavefronts Shared"]] or 0)
    cur["wave_x"] += float(r[ix["L1 Wavefronts Shared Excessive"]] or 0)
    for s in ("stall_barrier", "stall_short_sb", "stall_long_sb", "stall_math", "stall_mio", "stall_wait", "stall_not_selected", "stall_selected", "stall_dispatch"):
        cur[s] += float(r[ix[s]] or 0)
    if c == "bar":
        phases.append(cur); cur = collections.defaultdict(float); cur["_start"] = n + 1
phases.append(cur)
tot = sum(p["samples"] for p in phases)
for k, p in enumerate(phases):
    if p["inst"] == 0: continue
    print("phase %2d @%4d: samples %5.1f%%  inst/unit %7.1f  fp64 %7.1f smem %6.1f gmem %5.1f  waves/unit %7.1f (excess %6.1f)" % (
        k, p["_start"], 100 * p["samples"] / tot, p["inst"] / units, p["i:fp64"] / units, p["i:smem"] / units, p["i:gmem"] / units, p["wave"] / units, p["wave_x"] / units))
    st = {s: p[s] for s in p if s.startswith("stall_")}
    print("      stalls: " + " ".join("%s %.0f%%" % (s[6:], 100 * v / max(p["samples"], 1)) for s, v in sorted(st.items(), key=lambda kv: -kv[1])[:6]))
    im = sorted(((k2[2:], v) for k2, v in p.items() if k2.startswith("i:") and not k2.startswith("i:fp64")), key=lambda kv: -kv[1])[:8]
    print("      non-fp64: " + " ".join("%s %.0f" % (a, b / units) for a, b in im))
