"""The C-ABI library loads on a CPU-only box, exports every symbol include/helm_b200.h declares, and
refuses to compute without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "helm_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(hb_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported():
    from helm_b200 import engine
    L = engine.lib()
    syms = declared_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(L, s), "libhelm_b200.so does not export %s" % s
    # the python binding knows every one of them too
    assert sorted(engine.EXPORTS) == syms


def test_gate_struct_layout():
    from helm_b200 import engine
    assert C.sizeof(engine.HbGate) == 32 and engine.GATE_DTYPE.itemsize == 32
    assert engine.GATE_DTYPE.fields["out"][1] == 20 and engine.GATE_DTYPE.fields["lut"][1] == 24
    assert C.sizeof(engine.HbParams) == 48
    # hb_sop { u8 kind, u8 n_in, i8 coef[4], i16 cst, u32 in[4], u32 out, u64 table }
    assert engine.SOP_DTYPE.itemsize == 40 and engine.SOP_DTYPE.fields["cst"][1] == 6
    assert engine.SOP_DTYPE.fields["out"][1] == 24 and engine.SOP_DTYPE.fields["table"][1] == 32


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import helm_b200
    from oracle import oracle as O
    with pytest.raises(helm_b200.EngineError, match="no CUDA device"):
        helm_b200.Engine(O.PARAM_SETS["toy32"], device=0)


def test_product_never_imports_oracle():
    """oracle/ is test infrastructure: nothing under helm_b200/ may reference it."""
    for dp, _, fs in os.walk(os.path.join(ROOT, "helm_b200")):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp", "Makefile")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "tfhe_oracle" not in txt and "liboracle" not in txt and "from oracle" not in txt \
                    and "import oracle" not in txt, "%s references the oracle" % f


def test_integration_doc_declares_every_export():
    """INTEGRATION.md's Rust shim must bind every export of the header (15 of 25 were declared)."""
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    rust = set(re.findall(r"pub fn (hb_[a-z0-9_]+)\(", doc))
    assert rust == set(declared_symbols())
    assert "pub struct HbSop" in doc and "pub struct HbStats" in doc
