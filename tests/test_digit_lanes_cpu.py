"""The u32 gadget decomposition used by pbs16_kernel (helm_b200/csrc/digits_u32.cuh), compiled for the host: the scalar
closed form and the four-values-per-word lane form must both equal the tfhe-rs SignedDecomposer recurrence
(closest representable, then res/carry per level -- the rule gates.rs:255-274 reaches through tfhe-rs's blind rotation)
for every BL*L-bit state, in every lane."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_digit_lanes_match_recurrence(tmp_path):
    exe = str(tmp_path / "digit_lanes_check")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-I" + os.path.join(ROOT, "helm_b200", "csrc"), "-o", exe,
                           os.path.join(ROOT, "tests", "tools", "digit_lanes_check.cpp")])
    out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and out.stdout.startswith("ok "), out.stdout + out.stderr
