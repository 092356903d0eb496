#!/usr/bin/env python
"""Emits the Rust `extern "C"` block of INTEGRATION.md section 1 from include/helm_b200.h, so the shim's declarations cover every
export of the header (tests/test_abi.py checks INTEGRATION.md against the header).  usage: gen_rust_decls.py > decls.rs"""
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TY = {"hb_status": "i32", "int32_t": "i32", "uint32_t": "u32", "uint64_t": "u64", "size_t": "usize", "double": "f64",
      "void": "c_void", "char": "c_char", "hb_ctx": "HbCtx", "hb_wires": "HbWires", "hb_params": "HbParams", "hb_gate": "HbGate",
      "hb_sop": "HbSop", "hb_stats_t": "HbStats", "hb_graph": "HbGraph"}


def rust_type(c):
    c = c.strip()
    const = c.startswith("const ")
    c = c[6:] if const else c
    stars = c.count("*")
    base = TY[c.replace("*", "").strip()]
    for _ in range(stars):
        base = ("*const " if const else "*mut ") + base
    return base


def decls():
    src = open(os.path.join(ROOT, "include", "helm_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    out = []
    for m in re.finditer(r"^([a-z_0-9 ]+?[\s\*]+)(hb_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", src, flags=re.M | re.S):
        ret, name, args = m.group(1).strip(), m.group(2), " ".join(m.group(3).split())
        params = []
        for i, a in enumerate([x.strip() for x in args.split(",") if x.strip() and x.strip() != "void"]):
            mm = re.match(r"(.*?)([A-Za-z_][A-Za-z_0-9]*)?$", a)
            ty, nm = mm.group(1).strip(), mm.group(2)
            if not ty:                                   # unnamed parameter: the whole text is the type
                ty, nm = a, None
            if nm in TY and not ty.replace("const", "").strip():
                ty, nm = a, None
            params.append("%s: %s" % (nm or "a%d" % i, rust_type(ty)))
        r = rust_type(ret)
        out.append("    pub fn %s(%s)%s;" % (name, ", ".join(params), "" if r == "c_void" else " -> " + r))
    return out


if __name__ == "__main__":
    sys.stdout.write("extern \"C\" {\n" + "\n".join(decls()) + "\n}\n")
