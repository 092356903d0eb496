"""The register-level transforms of the PBS kernels (helm_b200/csrc/fft_regs.cuh) compiled for the host: the DIF form, the
FMA-form decimation-in-time butterflies and the variant that carries the negacyclic twist into its first stage must all
equal a direct long-double DFT (R = 16, 32; both signs) on digit-sized integer inputs."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_dft_forms_match_direct_dft(tmp_path):
    exe = str(tmp_path / "dft_forms_check")
    subprocess.check_call(["g++", "-O1", "-std=c++17", "-ffp-contract=off",
                           "-I" + os.path.join(ROOT, "tests", "tools", "stub"), "-I" + os.path.join(ROOT, "helm_b200", "csrc"),
                           "-o", exe, os.path.join(ROOT, "tests", "tools", "dft_forms_check.cpp")])
    out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and out.stdout.startswith("ok "), out.stdout + out.stderr
