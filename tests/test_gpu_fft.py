"""GPU parity of the hand-written R2C / C2R transforms (replacing cuFFT, ffter.hpp:46-78) against
numpy's double-precision FFT, against cuFFT through the reference harness, and -- at the sizes
the oracle cannot finish quickly -- through size-independent properties (Parseval, linearity,
round trip, a known tone)."""
import ctypes as C

import numpy as np
import pytest

import peasoup_b200 as pb
from peasoup_b200 import check, ptr

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

# tolerance: spectra within 1e-4 relative (north_star); measured ~3e-7 rms, ~3e-6 max
TOL_MAX, TOL_RMS = 2e-5, 2e-6


def relerr(got, want):
    rms = np.sqrt((np.abs(want) ** 2).mean())
    return np.abs(got - want).max() / rms, np.sqrt((np.abs(got - want) ** 2).mean()) / rms


@pytest.mark.parametrize("lg", [17, 18, 19, 20, 21, 22, 23, 24, 25])
def test_r2c_c2r_vs_numpy(lib, lg):
    n = 1 << lg
    rng = np.random.default_rng(lg)
    x = (rng.standard_normal(n) * 16).astype(np.float32)
    plan = pb.FFTPlan(n)
    dx = torch.from_numpy(x).cuda()
    dX = torch.zeros(2 * (n // 2 + 1), dtype=torch.float32, device="cuda")
    plan.r2c(dx, dX)
    X = dX.cpu().numpy().view(np.complex64)
    want = np.fft.rfft(x.astype(np.float64))
    emax, erms = relerr(X, want)
    assert emax < TOL_MAX and erms < TOL_RMS, (emax, erms)
    assert X[0].imag == 0 and X[-1].imag == 0
    # u8-like input with a DC term 1e4 x the noise bins: any float32 FFT spreads ~eps*|DC| along the
    # DC butterfly paths, so only the rms and a max bound of 1e-3 (relative to the noise bins) or four
    # float32 roundings of the DC term, whichever is larger (the DC term grows as n, the noise bins
    # as sqrt(n): at 2^25 samples the second bound takes over), hold
    xo = (x + 128).astype(np.float32)
    plan.r2c(torch.from_numpy(xo).cuda(), dX)
    Xo = dX.cpu().numpy().view(np.complex64)
    wo = np.fft.rfft(xo.astype(np.float64))
    nrms = np.sqrt((np.abs(wo[1:]) ** 2).mean())
    assert np.abs(Xo[1:] - wo[1:]).max() < max(1e-3 * nrms, 4 * 6e-8 * abs(wo[0]))
    assert np.sqrt((np.abs(Xo[1:] - wo[1:]) ** 2).mean()) / nrms < TOL_RMS
    assert abs(Xo[0] - wo[0]) / abs(wo[0]) < 2e-7
    # C2R on an exactly Hermitian spectrum: unnormalised (x n) like cufftExecC2R
    Xc = want.astype(np.complex64)
    dXc = torch.from_numpy(Xc.view(np.float32).copy()).cuda()
    dy = torch.zeros(n, dtype=torch.float32, device="cuda")
    plan.c2r(dXc, dy)
    want_y = np.fft.irfft(Xc.astype(np.complex128), n) * n
    emax, erms = relerr(dy.cpu().numpy(), want_y)
    assert emax < TOL_MAX and erms < TOL_RMS, (emax, erms)
    plan.close()


@pytest.mark.parametrize("lg", [17, 20, 23])
def test_vs_cufft_reference(lib, ref, lg):
    n = 1 << lg
    x = (np.random.default_rng(lg + 100).standard_normal(n)).astype(np.float32)
    dx = torch.from_numpy(x).cuda()
    a = torch.zeros(2 * (n // 2 + 1), dtype=torch.float32, device="cuda")
    b = torch.zeros_like(a)
    plan = pb.FFTPlan(n)
    plan.r2c(dx, a)
    ref.ref_r2c.argtypes = [C.c_uint, C.c_void_p, C.c_void_p]
    ref.ref_c2r.argtypes = [C.c_uint, C.c_void_p, C.c_void_p]
    assert ref.ref_r2c(n, ptr(dx), ptr(b)) == 0, ref.ref_last_error()
    torch.cuda.synchronize()
    emax, erms = relerr(a.cpu().numpy().view(np.complex64), b.cpu().numpy().view(np.complex64))
    assert emax < TOL_MAX and erms < TOL_RMS, (emax, erms)
    ya = torch.zeros(n, dtype=torch.float32, device="cuda")
    yb = torch.zeros_like(ya)
    b2 = b.clone()
    plan.c2r(b, ya)
    assert ref.ref_c2r(n, ptr(b2), ptr(yb)) == 0, ref.ref_last_error()
    torch.cuda.synchronize()
    emax, erms = relerr(ya.cpu().numpy(), yb.cpu().numpy())
    assert emax < TOL_MAX and erms < TOL_RMS, (emax, erms)
    plan.close()


@pytest.mark.parametrize("lg", [23, 24, 25])
def test_large_size_properties(lib, lg):
    n = 1 << lg
    plan = pb.FFTPlan(n)
    g = torch.Generator(device="cuda").manual_seed(lg)
    x = torch.randn(n, device="cuda", generator=g)
    y = torch.randn(n, device="cuda", generator=g)
    k0 = 1234567 % (n // 2)
    t = torch.arange(n, device="cuda", dtype=torch.float64)
    tone = torch.cos(2 * np.pi * k0 * t / n).float()
    X = torch.zeros(2 * (n // 2 + 1), dtype=torch.float32, device="cuda")
    Y = torch.zeros_like(X)
    S = torch.zeros_like(X)
    plan.r2c(x, X)
    plan.r2c(y, Y)
    plan.r2c(x + 2 * y, S)
    torch.cuda.synchronize()
    # linearity
    d = (S - (X + 2 * Y)).double()
    assert d.pow(2).mean().sqrt() / S.double().pow(2).mean().sqrt() < 1e-6
    # Parseval: sum x^2 = (|X0|^2 + |XM|^2 + 2 sum_{0<k<M} |Xk|^2) / n
    Xc = torch.view_as_complex(X.view(-1, 2)).to(torch.complex128)
    p = Xc.abs().pow(2)
    lhs = x.double().pow(2).sum()
    rhs = (p[0] + p[-1] + 2 * p[1:-1].sum()) / n
    assert abs(lhs - rhs) / lhs < 1e-6
    # a pure tone lands in one bin with amplitude n/2
    T = torch.zeros_like(X)
    plan.r2c(tone, T)
    Tc = torch.view_as_complex(T.view(-1, 2))
    assert int(Tc.abs().argmax()) == k0
    assert abs(float(Tc.abs()[k0]) - n / 2) / (n / 2) < 1e-5
    # round trip: c2r(r2c(x)) = n x
    back = torch.zeros(n, dtype=torch.float32, device="cuda")
    plan.c2r(X, back)
    err = (back.double() / n - x.double()).abs().max()
    assert err < 2e-5
    plan.close()


def test_plan_rejects_unsupported_sizes(lib):
    for n in (0, 1, (1 << 24) + 1, 1 << 26):
        h = C.c_void_p()
        assert lib.ps_fft_plan_create(n, C.byref(h)) == -2
        assert b"power of two" in lib.ps_last_error()


@pytest.mark.parametrize("n", [1000, 1 << 16, 99991, (1 << 17) + 2, 187520, 3 << 16, 1000001])
def test_arbitrary_length_vs_numpy(lib, n):
    """Lengths cuFFT would take but the power-of-two engine does not (the coincidencer transforms
    the whole observation, coincidencer.cpp:137-139): chirp-z on top of the 2^k transforms."""
    rng = np.random.default_rng(n)
    x = (rng.standard_normal(n) * 16).astype(np.float32)
    plan = pb.FFTPlan(n)
    dX = torch.zeros(2 * (n // 2 + 1), dtype=torch.float32, device="cuda")
    plan.r2c(torch.from_numpy(x).cuda(), dX)
    X = dX.cpu().numpy().view(np.complex64)
    want = np.fft.rfft(x.astype(np.float64))
    emax, erms = relerr(X, want)
    assert emax < 5e-5 and erms < 5e-6, (emax, erms)
    # u8-like input (large DC): the mean is taken out before the convolution and put back into
    # bin 0, so the other bins are as accurate as for zero-mean input
    xo = (x + 128).astype(np.float32)
    plan.r2c(torch.from_numpy(xo).cuda(), dX)
    Xo = dX.cpu().numpy().view(np.complex64)
    wo = np.fft.rfft(xo.astype(np.float64))
    nrms = np.sqrt((np.abs(wo[1:]) ** 2).mean())
    assert np.sqrt((np.abs(Xo[1:] - wo[1:]) ** 2).mean()) / nrms < 5e-6
    assert np.abs(Xo[1:] - wo[1:]).max() / nrms < 5e-5
    assert abs(Xo[0] - wo[0]) / abs(wo[0]) < 1e-6
    # C2R, unnormalised
    Xc = want.astype(np.complex64)
    dy = torch.zeros(n, dtype=torch.float32, device="cuda")
    plan.c2r(torch.from_numpy(Xc.view(np.float32).copy()).cuda(), dy)
    want_y = np.fft.irfft(Xc.astype(np.complex128), n) * n
    emax, erms = relerr(dy.cpu().numpy(), want_y)
    assert emax < 5e-5 and erms < 5e-6, (emax, erms)
    plan.close()
