"""The butterflies of csrc/mel_fft.cuh compiled for the host with g++ and checked against numpy:
the 400-point real FFT the mel kernel runs (25 x rdft16 -> twiddle -> 9 x dft25)."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    out = tmp_path_factory.mktemp("fft") / "libfft_host.so"
    src = os.path.join(ROOT, "tests", "fft_host_harness.cpp")
    inc = os.path.join(ROOT, "qwen3-asr.cpp_b200", "csrc")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-I", inc, "-o", str(out), src])
    lib = ctypes.CDLL(str(out))
    lib.fft400_power.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    return lib


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_fft400_power_matches_numpy(harness, seed):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal(400).astype(np.float32)
    p = np.zeros(201, np.float32)
    harness.fft400_power(x.ctypes.data, p.ctypes.data)
    want = np.abs(np.fft.rfft(x.astype(np.float64))) ** 2
    assert np.abs(p - want).max() <= 2e-5 * want.max()


def test_fft400_impulse_and_tone(harness):
    x = np.zeros(400, np.float32); x[0] = 1.0
    p = np.zeros(201, np.float32)
    harness.fft400_power(x.ctypes.data, p.ctypes.data)
    assert np.allclose(p, 1.0, atol=1e-6)
    x = np.cos(2 * np.pi * 37 * np.arange(400) / 400).astype(np.float32)
    harness.fft400_power(x.ctypes.data, p.ctypes.data)
    assert abs(p[37] - 200.0 ** 2) < 1.0 and np.delete(p, 37).max() < 1e-3
