"""GPU parity of every staged C-ABI entry point (include/peasoup_b200.h) against the CPU oracle
and -- when oracle/_ref was built -- against the UNMODIFIED reference kernels on the same device
buffers.  Integer / index / selection work is bit-exact; float element-wise work is bit-exact
where the arithmetic is IEEE (and within 2 ulp where the reference uses MUFU approximations)."""
import ctypes as C

import numpy as np
import pytest

import peasoup_b200 as pb
from peasoup_b200 import check, ptr

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def rcheck(ref, rc):
    assert rc == 0, ref.ref_last_error()


# ------------------------------------------------------------------ resampling (bit-exact)
@pytest.mark.parametrize("n,a,tsamp", [(1 << 17, 5.0, 0.00032), (1 << 17, -5.0, 0.00032), (4194304, 125.5, 64e-6),
                                       (4194304, -250.0, 64e-6), (1 << 23, 50.0, 64e-6), (1 << 23, -31.4, 64e-6),
                                       (1 << 24, 500.0, 64e-6), (1000003, 77.7, 1e-4), (1 << 20, 0.0, 64e-6)])
@pytest.mark.parametrize("which", [1, 2])
def test_resample_index_maps_bit_exact(lib, oracle, n, a, tsamp, which):
    d_idx = torch.zeros(n, dtype=torch.int32, device="cuda")
    check(lib.ps_resample_index_map(ptr(d_idx), n, a, tsamp, which, None))
    got = d_idx.cpu().numpy().view(np.uint32).astype(np.uint64)
    want = oracle.resample_index(n, a, tsamp, which)
    assert np.array_equal(got, want)
    assert got.max() < n
    if a > 0 and which == 2:
        assert np.all(got <= np.arange(n, dtype=np.uint64))  # a>0 reads behind


def test_resampling_test_pattern_vs_reference(lib, oracle, ref):
    # src/resampling_test.cpp:14-64: N=4194304, tsamp=64us, acc=125.5, pattern i%451
    n, a, tsamp = 4194304, 125.5, 64e-6
    x = dev((np.arange(n) % 451).astype(np.float32))
    for name_ps, name_ref in (("ps_resampleII", "ref_resampleII"), ("ps_resample", "ref_resample")):
        y = torch.zeros_like(x)
        z = torch.zeros_like(x)
        check(getattr(lib, name_ps)(ptr(x), ptr(y), n, a, tsamp, None))
        fn = getattr(ref, name_ref)
        fn.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_float, C.c_float]
        rcheck(ref, fn(ptr(x), ptr(z), n, a, tsamp))
        torch.cuda.synchronize()
        assert torch.equal(y, z)
    assert np.array_equal(y.cpu().numpy(), oracle.resample(x.cpu().numpy(), a, tsamp, 1))


def test_resample_empty(lib):
    check(lib.ps_resampleII(None, None, 0, 1.0, 1.0, None))


# ------------------------------------------------------------------ spectra / normalise
@pytest.mark.parametrize("nbins", [1, 7, 65537, (1 << 22) + 1])
def test_form_power_series(lib, oracle, nbins, request):
    rng = np.random.default_rng(nbins)
    X = (rng.standard_normal(nbins) * 100 + 1j * rng.standard_normal(nbins) * 100).astype(np.complex64)
    dX = dev(X.view(np.float32))
    out = torch.zeros(nbins, dtype=torch.float32, device="cuda")
    check(lib.ps_form_power_series(ptr(dX), ptr(out), nbins, 1, None))
    assert np.array_equal(out.cpu().numpy(), oracle.form_power(X, 1))  # IEEE ops only: bit-exact
    check(lib.ps_form_power_series(ptr(dX), ptr(out), nbins, 0, None))
    want = oracle.form_power(X, 0)
    assert np.allclose(out.cpu().numpy(), want, rtol=3e-7, atol=0)  # rsqrtf is a 2-ulp approximation


def test_form_power_series_vs_reference(lib, ref):
    nbins = 1 << 20
    rng = np.random.default_rng(5)
    X = (rng.standard_normal(nbins) + 1j * rng.standard_normal(nbins)).astype(np.complex64)
    dX = dev(X.view(np.float32))
    ref.ref_form_power_series.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
    for way in (0, 1):
        a = torch.zeros(nbins, dtype=torch.float32, device="cuda")
        b = torch.zeros_like(a)
        check(lib.ps_form_power_series(ptr(dX), ptr(a), nbins, way, None))
        rcheck(ref, ref.ref_form_power_series(ptr(dX), ptr(b), nbins, way))
        torch.cuda.synchronize()
        assert torch.equal(a, b), way


def test_power_is_nan_at_zero_like_reference(lib):
    dX = torch.zeros(8, dtype=torch.float32, device="cuda")
    out = torch.zeros(4, dtype=torch.float32, device="cuda")
    check(lib.ps_form_power_series(ptr(dX), ptr(out), 4, 0, None))
    assert torch.isnan(out).all()  # z*rsqrtf(z) at z == 0 (kernels.cu:223-224)


def test_normalise(lib, oracle, ref):
    n = (1 << 20) + 3
    x = np.random.default_rng(1).standard_normal(n).astype(np.float32) * 1e6 + 5e6
    d = dev(x)
    check(lib.ps_normalise(ptr(d), 4.9e6, 1.1e6, n, None))
    assert np.array_equal(d.cpu().numpy(), oracle.normalise(x, 4.9e6, 1.1e6))
    d2 = dev(x)
    ref.ref_normalise.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_uint]
    rcheck(ref, ref.ref_normalise(ptr(d2), 4.9e6, 1.1e6, n))
    assert torch.equal(d, d2)


# ------------------------------------------------------------------ harmonic sums
def harmonic_sum_gpu(lib, x, nfolds):
    n = len(x)
    d = dev(x)
    out = torch.zeros((max(nfolds, 1), n), dtype=torch.float32, device="cuda")
    ptrs = (C.c_void_p * 5)(*[out[i].data_ptr() if i < nfolds else None for i in range(5)])
    check(lib.ps_harmonic_sum(ptr(d), ptrs, n, nfolds, None))
    return d, out


@pytest.mark.parametrize("n,nfolds", [(1, 4), (31, 5), (65537, 4), (10_000_000, 4), ((1 << 22) + 1, 5), (1000, 0)])
def test_harmonic_sum_vs_oracle(lib, oracle, n, nfolds):
    if n == 10_000_000:  # src/harmonic_sum_test.cpp:13-55 pattern: 1 every 32 bins
        x = np.zeros(n, np.float32)
        x[::32] = 1
    else:
        x = np.random.default_rng(n).standard_normal(n).astype(np.float32)
    _, out = harmonic_sum_gpu(lib, x, nfolds)
    if nfolds:
        want = oracle.harmonic_sum(x, nfolds)
        got = out.cpu().numpy()[:nfolds]
        assert np.allclose(got, want, rtol=2e-7, atol=1e-7)  # double-precision rsqrt constant: <= 1 ulp
        assert np.array_equal(got[1::2], want[1::2]) or nfolds < 2  # folds 2 and 4 scale by 0.5/0.25: exact


def test_harmonic_sum_vs_reference_bit_exact(lib, ref):
    n, nfolds = (1 << 22) + 1, 5
    x = np.random.default_rng(3).standard_normal(n).astype(np.float32)
    d, out = harmonic_sum_gpu(lib, x, nfolds)
    want = torch.zeros_like(out)
    ref.ref_harmonic_sum.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint]
    rcheck(ref, ref.ref_harmonic_sum(ptr(d), ptr(want), n, nfolds))
    torch.cuda.synchronize()
    assert torch.equal(out, want)


def test_harmonic_sum_rejects_six_folds(lib):
    ptrs = (C.c_void_p * 5)()
    assert lib.ps_harmonic_sum(None, ptrs, 10, 6, None) == -2
    assert b"5 folds" in lib.ps_last_error()


# ------------------------------------------------------------------ peaks
@pytest.mark.parametrize("n,start,thresh", [(100000, 0, 3.0), (4194305, 107, 4.0), (65537, 65530, -10.0), (1000, 2000, 0.0),
                                            (50, 0, 100.0)])
def test_find_peaks(lib, oracle, n, start, thresh):
    x = np.random.default_rng(n).standard_normal(n).astype(np.float32)
    d = dev(x)
    ws = C.c_void_p()
    check(lib.ps_peak_workspace_create(200000, C.byref(ws)))
    hi = np.zeros(200000, np.int32)
    hs = np.zeros(200000, np.float32)
    c = check(lib.ps_find_peaks(n, start, ptr(d), thresh, ptr(hi), ptr(hs), len(hi), ws, None))
    wi, wsn = oracle.find_peaks(x, n, start, thresh)
    assert c == len(wi)
    assert np.array_equal(hi[:c], wi) and np.array_equal(hs[:c], wsn)
    lib.ps_peak_workspace_destroy(ws)


def test_find_peaks_overflow_is_an_error(lib):
    x = dev(np.ones(5000, np.float32))
    ws = C.c_void_p()
    check(lib.ps_peak_workspace_create(100, C.byref(ws)))
    hi = np.zeros(100, np.int32)
    hs = np.zeros(100, np.float32)
    assert lib.ps_find_peaks(5000, 10, ptr(x), 0.5, ptr(hi), ptr(hs), 100, ws, None) == -3
    # ... and the workspace says how many there were, so the caller can grow and search again
    # (PeakFinder::find_candidates of the mirror does: 1.5x the count, like the fused path's peak list)
    need = lib.ps_peak_workspace_needed(ws)
    assert need == 4990
    lib.ps_peak_workspace_destroy(ws)
    check(lib.ps_peak_workspace_create(need + need // 2, C.byref(ws)))
    hi = np.zeros(need + need // 2, np.int32)
    hs = np.zeros(need + need // 2, np.float32)
    assert lib.ps_find_peaks(5000, 10, ptr(x), 0.5, ptr(hi), ptr(hs), len(hi), ws, None) == need
    assert np.array_equal(hi[:need], np.arange(10, 5000))
    lib.ps_peak_workspace_destroy(ws)


def test_find_peaks_vs_reference(lib, ref):
    n = 1 << 20
    x = np.random.default_rng(9).standard_normal(n).astype(np.float32)
    d = dev(x)
    ws = C.c_void_p()
    check(lib.ps_peak_workspace_create(100000, C.byref(ws)))
    hi, hs = np.zeros(100000, np.int32), np.zeros(100000, np.float32)
    c = check(lib.ps_find_peaks(n, 33, ptr(d), 3.5, ptr(hi), ptr(hs), len(hi), ws, None))
    ri, rs = np.zeros(n, np.int32), np.zeros(n, np.float32)
    cnt = C.c_int()
    ref.ref_find_peaks.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.POINTER(C.c_int)]
    rcheck(ref, ref.ref_find_peaks(n, 33, ptr(d), 3.5, ptr(ri), ptr(rs), C.byref(cnt)))
    assert c == cnt.value and np.array_equal(hi[:c], ri[:c]) and np.array_equal(hs[:c], rs[:c])
    lib.ps_peak_workspace_destroy(ws)


# ------------------------------------------------------------------ rednoise
@pytest.mark.parametrize("count", [1, 2, 3, 4, 5, 9, 10, 65537, 4194305])
def test_median_scrunch5(lib, oracle, count):
    x = np.random.default_rng(count).random(count).astype(np.float32)
    x[::7] = x[0]  # ties
    d = dev(x)
    out = torch.zeros(max(1, count // 5), dtype=torch.float32, device="cuda")
    check(lib.ps_median_scrunch5(ptr(d), count, ptr(out), None))
    assert np.array_equal(out.cpu().numpy(), oracle.median_scrunch5(x))


@pytest.mark.parametrize("nin,nout", [(13107, 65537), (2621, 65537), (838861, 4194305), (2, 10), (5, 5)])
def test_linear_stretch(lib, oracle, nin, nout):
    x = np.random.default_rng(nin).random(nin).astype(np.float32)
    d = dev(x)
    out = torch.zeros(nout, dtype=torch.float32, device="cuda")
    check(lib.ps_linear_stretch(ptr(d), nin, ptr(out), nout, None))
    assert np.array_equal(out.cpu().numpy(), oracle.linear_stretch(x, nout))


@pytest.mark.parametrize("size,tsamp", [(1 << 17, 0.00032), (1 << 20, 64e-6), (1 << 23, 64e-6)])
def test_running_median_vs_oracle_and_reference(lib, oracle, size, tsamp, request):
    cfg = oracle.SearchConfig(size, tsamp)
    nb = cfg.nbins
    rng = np.random.default_rng(size)
    p = (np.abs(rng.standard_normal(nb)) * (1 + 50.0 / (1 + np.arange(nb) / 50.0))).astype(np.float32)
    d = dev(p)
    med = torch.zeros(nb, dtype=torch.float32, device="cuda")
    work = torch.zeros(nb // 5 + nb // 25 + nb // 125 + 8, dtype=torch.float32, device="cuda")
    check(lib.ps_running_median(ptr(d), nb, cfg.bin_width, 0.05, 0.5, ptr(med), ptr(work), None))
    want = oracle.running_median(p, cfg.bin_width)
    assert np.array_equal(med.cpu().numpy(), want)
    try:
        ref = request.getfixturevalue("ref")
    except pytest.skip.Exception:
        return
    rmed = torch.zeros_like(med)
    ref.ref_running_median.argtypes = [C.c_void_p, C.c_uint, C.c_double, C.c_void_p, C.c_void_p]
    rcheck(ref, ref.ref_running_median(ptr(d), nb, cfg.bin_width, ptr(rmed), None))
    g, r = med.cpu().numpy(), rmed.cpu().numpy()
    # the reference reads one element past its 125-bin level at the very last bin (kernels.cu:994)
    assert np.array_equal(g[:-1], r[:-1])


def test_running_median_rejects_short_input(lib):
    d = torch.zeros(100, dtype=torch.float32, device="cuda")
    assert lib.ps_running_median(ptr(d), 100, 0.001, 0.05, 0.5, ptr(d), ptr(d), None) == -2
    assert b"running_median" in lib.ps_last_error()


def test_divide_and_zap(lib, oracle, request):
    cfg = oracle.SearchConfig(1 << 20, 64e-6)
    nb = cfg.nbins
    rng = np.random.default_rng(2)
    X = (rng.standard_normal(nb) * 1e3 + 1j * rng.standard_normal(nb) * 1e3).astype(np.complex64)
    f = (np.abs(rng.standard_normal(nb)) * 1e3 + 1).astype(np.float32)
    dX = dev(X.view(np.float32))
    df = dev(f)
    check(lib.ps_divide_c_by_f(ptr(dX), ptr(df), nb, None))
    want = oracle.divide_c_by_f(X, f)
    assert np.array_equal(dX.cpu().numpy().view(np.complex64), want)
    assert np.all(want[:5] == 0)
    birdies = np.array([50.0, 100.0, 150.0, 0.01, 7812.4, 9000.0], np.float32)
    widths = np.array([0.1, 0.15, 0.15, 0.5, 0.2, 0.1], np.float32)
    db, dw = dev(birdies), dev(widths)
    check(lib.ps_zap_birdies(ptr(dX), ptr(db), ptr(dw), cfg.bin_width_f, len(birdies), nb, None))
    want2 = oracle.zap(want, birdies, widths, cfg.bin_width_f)
    got = dX.cpu().numpy().view(np.complex64)
    assert np.array_equal(got, want2)
    assert (got == 1).sum() > 100
    try:
        ref = request.getfixturevalue("ref")
    except pytest.skip.Exception:
        return
    rX = dev(X.view(np.float32))
    ref.ref_divide_c_by_f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint]
    ref.ref_zap_birdies.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_uint, C.c_uint]
    rcheck(ref, ref.ref_divide_c_by_f(ptr(rX), ptr(df), nb))
    rcheck(ref, ref.ref_zap_birdies(ptr(rX), ptr(db), ptr(dw), cfg.bin_width_f, len(birdies), nb))
    assert torch.equal(rX, dX)


@pytest.mark.parametrize("size", [1 << 20, 1 << 23])
def test_zap_47tuc_zaplist_vs_reference(lib, oracle, ref, size):
    """The reference's survey zaplist misc/47tuc.zaplist (104 birdies, kept as tests/golden/47tuc.zaplist)
    through ps_zap_birdies against zap_birdies_kernel of the unmodified reference build
    (kernels.cu:1036-1069) and the oracle: bit-exact, including birdies whose range is clipped at the
    spectrum's ends."""
    import os
    from conftest import GOLDEN
    zl = np.loadtxt(os.path.join(GOLDEN, "47tuc.zaplist"), dtype=np.float32)
    assert zl.shape == (104, 2)
    cfg = oracle.SearchConfig(size, 64e-6)
    nb = cfg.nbins
    rng = np.random.default_rng(47)
    X = (rng.standard_normal(nb) + 1j * rng.standard_normal(nb)).astype(np.complex64)
    dX, rX = dev(X.view(np.float32)), dev(X.view(np.float32))
    db, dw = dev(zl[:, 0]), dev(zl[:, 1])
    check(lib.ps_zap_birdies(ptr(dX), ptr(db), ptr(dw), cfg.bin_width_f, 104, nb, None))
    ref.ref_zap_birdies.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_uint, C.c_uint]
    rcheck(ref, ref.ref_zap_birdies(ptr(rX), ptr(db), ptr(dw), cfg.bin_width_f, 104, nb))
    torch.cuda.synchronize()
    assert torch.equal(dX, rX)
    got = dX.cpu().numpy().view(np.complex64)
    assert np.array_equal(got, oracle.zap(X, zl[:, 0], zl[:, 1], cfg.bin_width_f))
    assert (got == 1).sum() >= 104


# ------------------------------------------------------------------ stats / fill / convert
def test_stats_fill_convert(lib, oracle, request):
    n = 4194305
    x = (np.abs(np.random.default_rng(4).standard_normal(n)) * 3 + 1).astype(np.float32)
    d = dev(x)
    m, r = C.c_float(), C.c_float()
    check(lib.ps_mean(ptr(d), n, 0, C.byref(m)))
    check(lib.ps_rms(ptr(d), n, 0, C.byref(r)))
    wm, wr, ws = oracle.stats(x)
    # both accumulate in double (different association) and narrow identically: <= 1 ulp
    assert abs(m.value - wm) <= 1.2e-7 * wm and abs(r.value - wr) <= 1.2e-7 * wr
    check(lib.ps_mean(ptr(d), n, 1000, C.byref(m)))
    assert abs(m.value - x[1000:].astype(np.float64).mean()) < 1e-6 * m.value
    try:
        ref = request.getfixturevalue("ref")
        rm, rr, rs = C.c_float(), C.c_float(), C.c_float()
        ref.ref_stats.argtypes = [C.c_void_p, C.c_uint, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_float)]
        rcheck(ref, ref.ref_stats(ptr(d), n, C.byref(rm), C.byref(rr), C.byref(rs)))
        # Thrust float tree reductions: order unpinned -> tolerance
        assert abs(rm.value - wm) < 2e-6 * wm and abs(rr.value - wr) < 2e-6 * wr
    except pytest.skip.Exception:
        pass
    check(lib.ps_fill(ptr(d) + 4 * 10, ptr(d) + 4 * 1000, 7.5, None))
    h = d.cpu().numpy()
    assert np.all(h[10:1000] == 7.5) and h[9] == x[9] and h[1000] == x[1000]
    b = np.random.default_rng(0).integers(0, 256, 100003, dtype=np.uint8)
    db = dev(b)
    out = torch.zeros(len(b), dtype=torch.float32, device="cuda")
    check(lib.ps_convert_u8_f32(ptr(db), ptr(out), len(b), None))
    assert np.array_equal(out.cpu().numpy(), b.astype(np.float32))
