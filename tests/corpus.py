"""The hdl-benchmarks netlist corpus as committed fixtures (tests/golden/yosys, imported by scripts/import_fixtures.py from
/root/reference/hdl-benchmarks/netlists @ d949d36): which files exist, how each is run."""
import gzip
import os

YOSYS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "yosys")
# behavioural designs (arithmetic mode sources) and the preprocessor's unit fixture are not gate / LUT netlists
NOT_NETLISTS = {"chi_squared.v", "v2-euclidean-distance.v", "lut-multibit-input-test.v"}
# a DFF whose D pin is the literal 1'h1: the reference preprocessor writes `dff g(1'h1, q);` (hdl-benchmarks/src/lib.rs:754-812)
# and HELM's levelizer then panics on the unknown wire "1'h1" (src/circuit.rs:199-207) -- reference result undefined
REFERENCE_PANICS = {"s13207-lut-2-1.v", "s15850-lut-2-1.v"}


def names():
    out = []
    for f in sorted(os.listdir(YOSYS)):
        base = f[:-3] if f.endswith(".gz") else f
        if base.endswith(".v") and base not in NOT_NETLISTS:
            out.append(base)
    return out


def text(base):
    p = os.path.join(YOSYS, base)
    if os.path.exists(p):
        return open(p).read()
    with gzip.open(p + ".gz", "rt") as f:
        return f.read()
