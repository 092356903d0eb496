#!/usr/bin/env python
"""Copies the benchmark DATA files the reference's own tests and BASELINE.json configs use (structural-Verilog
netlists and input/golden CSVs from the hdl-benchmarks submodule -- circuit descriptions, not reference code) into
tests/golden/netlists and tests/golden/test-cases, because /root/reference does not exist on the GPU box.
Run here (CPU container) only.  Provenance: TrustworthyComputing/hdl-benchmarks @ d949d36."""
import os, shutil, sys
SRC = "/root/reference/hdl-benchmarks"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NETS = ["processed-netlists/" + f for f in ("s27.v", "2-bit-adder.v", "c17.v", "16-bit-mult-gates.v", "8-bit-adder-lut-2-1.v",
        "8-bit-adder-lut-3-1.v", "chi_squared_arith.v", "invalid.v", "gates.v", "c17_lut_2.v", "c17_lut_4.v", "c880_lut_4.v",
        "c7552_lut_4.v")] + \
       ["netlists/" + f for f in ("c880.v", "c7552.v", "c17.v", "c17_lut_4.v", "c880_lut_4.v", "c7552_lut_4.v",
        "lut-multibit-input-test.v", "8-bit-adder-lut.v", "c1355-lut-2-1.v", "c1355.v", "s27.v", "s386.v",
        "16-bit-mult-gates.v")] + ["designs/chi_squared.v"]
CASES = ["2-bit-adder.inputs.csv", "2-bit-adder.outputs.csv", "8-bit-adder.inputs.csv", "bool_array.input.csv",
         "16-bit-mult.inputs.csv"] + ["chi_squared_arith_%d.%s.csv" % (i, k) for i in range(1, 5) for k in ("inputs", "outputs")]
for rel in NETS:
    sub = "yosys" if rel.startswith(("netlists/", "designs/")) else "netlists"
    dst = os.path.join(ROOT, "tests", "golden", sub)
    os.makedirs(dst, exist_ok=True)
    shutil.copy(os.path.join(SRC, rel), dst)
dst = os.path.join(ROOT, "tests", "golden", "test-cases")
os.makedirs(dst, exist_ok=True)
for f in CASES:
    shutil.copy(os.path.join(SRC, "test-cases", f), dst)
# the large benchmark circuits of the reference's test-cases directory, gzip-compressed (1.7 MB of Yosys text each)
import gzip
BIG = ["netlists/aes_core_gates.v", "netlists/chi_squared_gates.v", "netlists/v2_euclidean_gates.v", "designs/v2-euclidean-distance.v"]
BIG_CASES = ["aes.inputs.csv", "chi_squared_bin_1.inputs.csv", "v2-euclidean-distance-bin.inputs.csv", "v2-euclidean-distance.inputs.csv",
             "32-bit-mult.inputs.csv"]
for rel in BIG:
    with open(os.path.join(SRC, rel), "rb") as f, gzip.GzipFile(os.path.join(ROOT, "tests", "golden", "yosys", os.path.basename(rel) + ".gz"), "wb", mtime=0) as g:
        g.write(f.read())
with open(os.path.join(SRC, "processed-netlists", "32-bit-mult-gates.v"), "rb") as f, \
        gzip.GzipFile(os.path.join(ROOT, "tests", "golden", "netlists", "32-bit-mult-gates.v.gz"), "wb", mtime=0) as g:
    g.write(f.read())
shutil.copy(os.path.join(SRC, "processed-netlists", "5x5_x_5x5-mmult-arith.v"), os.path.join(ROOT, "tests", "golden", "netlists"))
for f in BIG_CASES:
    shutil.copy(os.path.join(SRC, "test-cases", f), dst)
# the rest of the netlist corpus (north_star: "every HELM benchmark netlist"): every Yosys netlist of hdl-benchmarks/netlists not
# imported above, gzip-compressed; tests/test_gpu_corpus.py runs them all.  The 10 files of .MISSING_LARGE_BLOBS are absent upstream.
have = {os.path.basename(r) for r in NETS if r.startswith("netlists/")} | {os.path.basename(r) for r in BIG}
corpus = 0
for f in sorted(os.listdir(os.path.join(SRC, "netlists"))):
    if not f.endswith(".v") or f in have:
        continue
    with open(os.path.join(SRC, "netlists", f), "rb") as src, \
            gzip.GzipFile(os.path.join(ROOT, "tests", "golden", "yosys", f + ".gz"), "wb", mtime=0) as g:
        g.write(src.read())
    corpus += 1
shutil.copy(os.path.join(SRC, "test-cases", "128-bit-mult.inputs.csv"), dst)
print("imported", len(NETS) + len(BIG) + 2 + corpus, "netlists and", len(CASES) + len(BIG_CASES) + 1, "csv files")
