"""CPU execution of the FFT kernels' phase functions, thread by thread, against a double-precision
transform (tests/cpp/*.cpp compile the kernels' own __host__ __device__ code with g++).  This is
how index maps, twiddle tables and shared-memory layouts are checked without a GPU."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(tmp_path, name):
    exe = str(tmp_path / name)
    subprocess.run(["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "sdr-for-android-lib_b200", "csrc"),
                    "-o", exe, os.path.join(ROOT, "tests", "cpp", name + ".cpp")], check=True)
    return exe


def test_fft16k_phase_functions(tmp_path):
    r = subprocess.run([_build(tmp_path, "fft_sim")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr


@pytest.mark.parametrize("radix", [16, 32])
def test_four_step_phase_functions(tmp_path, radix):
    r = subprocess.run([_build(tmp_path, "fft_big_sim"), str(radix)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 bad bins" in r.stdout


def test_small_fft_phase_functions(tmp_path):
    """256 .. 8192 points through the one-CTA kernel's phase functions, 32768 points through the two-pass scheme
    built on it, and the bank-pair audit of its exchange buffer."""
    r = subprocess.run([_build(tmp_path, "fft_small_sim")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "all sizes ok" in r.stdout
