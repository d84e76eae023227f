"""Host execution of the CTA FFT phases (peasoup_b200/csrc/fft_core.cuh compiled as C++) against
numpy: checks digit maps, shared-memory layouts, twiddles and the four-step algebra without a GPU."""
import ctypes
import os
import subprocess

import numpy as np
import pytest
from conftest import REPO


@pytest.fixture(scope="module")
def emul():
    src = os.path.join(REPO, "tests", "cpu_emul", "fft_emul.cpp")
    so = os.path.join(REPO, "build", "libfft_emul.so")
    os.makedirs(os.path.dirname(so), exist_ok=True)
    hdr = os.path.join(REPO, "peasoup_b200", "csrc", "fft_core.cuh")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", so, src], check=True)
    return ctypes.CDLL(so)


def p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


@pytest.mark.parametrize("n1,n2", [(256, 256), (256, 512), (512, 1024), (1024, 1024), (1024, 2048), (2048, 2048),
                                   (2048, 4096), (4096, 4096)])
@pytest.mark.parametrize("inverse", [0, 1])
def test_complex_two_pass(emul, n1, n2, inverse):
    if n1 * n2 > (1 << 22) and inverse:
        pytest.skip("large inverse covered by the forward case")
    rng = np.random.default_rng(n1 + n2 + inverse)
    M = n1 * n2
    x = (rng.standard_normal(M) + 1j * rng.standard_normal(M)).astype(np.complex64)
    out = np.zeros(M, np.complex64)
    assert emul.emul_cfft(n1, n2, inverse, p(x), p(out)) == 0
    ref = np.fft.ifft(x.astype(np.complex128)) * M if inverse else np.fft.fft(x.astype(np.complex128))
    rms = np.sqrt((np.abs(ref) ** 2).mean())
    assert np.abs(out - ref).max() / rms < 2e-5
    assert np.sqrt((np.abs(out - ref) ** 2).mean()) / rms < 2e-6


@pytest.mark.parametrize("n", [1 << 17, 1 << 18, 1 << 20])
def test_real_transforms(emul, n):
    rng = np.random.default_rng(n)
    x = rng.standard_normal(n).astype(np.float32)
    X = np.zeros(n // 2 + 1, np.complex64)
    assert emul.emul_r2c(n, p(x), p(X)) == 0
    ref = np.fft.rfft(x.astype(np.float64))
    assert np.abs(X - ref).max() / np.sqrt((np.abs(ref) ** 2).mean()) < 2e-5
    y = np.zeros(n, np.float32)
    Xc = ref.astype(np.complex64)
    assert emul.emul_c2r(n, p(Xc), p(y)) == 0
    r2 = np.fft.irfft(Xc.astype(np.complex128), n) * n
    assert np.abs(y - r2).max() / np.sqrt((r2 ** 2).mean()) < 2e-5


def test_impulse_and_tone(emul):
    n1 = n2 = 256
    M = n1 * n2
    x = np.zeros(M, np.complex64)
    x[1] = 1
    out = np.zeros(M, np.complex64)
    emul.emul_cfft(n1, n2, 0, p(x), p(out))
    k = np.arange(M)
    assert np.abs(out - np.exp(-2j * np.pi * k / M)).max() < 1e-5
    x = np.exp(2j * np.pi * 12345 * k / M).astype(np.complex64)
    emul.emul_cfft(n1, n2, 0, p(x), p(out))
    assert np.argmax(np.abs(out)) == 12345
