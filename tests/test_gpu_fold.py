"""GPU parity of candidate folding and fold optimisation through the C ABI:
ps_fold_timeseries / ps_fold_optimise against oracle/fold.py on seeded inputs, and the whole
program with --npdmp against the reference's golden FOLD blocks, folded_snr and opt_period."""
import ctypes as C
import os

import numpy as np
import pytest

import peasoup_b200 as pb
from conftest import GOLDEN
from test_gpu_filterbank import run_tutorial
from test_oracle_golden import rows_matching

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def pulsar_series(n, tsamp, period, width=0.04, amp=3.0, seed=0):
    rng = np.random.default_rng(seed)
    t = np.arange(n) * tsamp
    phase = (t / period) % 1.0
    return (rng.normal(0, 1, n) + amp * (np.abs(phase - 0.37) < width / 2)).astype(np.float32)


def gpu_fold(lib, x, period, tsamp, nbins=64, nints=16):
    d_in = dev(x)
    d_out = torch.zeros(nints * nbins, dtype=torch.float32, device="cuda")
    pb.check(lib.ps_fold_timeseries(d_in.data_ptr(), d_out.data_ptr(), len(x), nints, float(period), float(tsamp), nbins, None))
    torch.cuda.synchronize()
    return d_out.cpu().numpy().reshape(nints, nbins)


@pytest.mark.parametrize("n,period,nbins,nints", [(1 << 17, 0.25, 64, 16), (1 << 17, 0.0123456, 64, 16), (100000, 0.0817, 64, 16),
                                                  (1 << 20, 1.2345, 128, 8), (5000, 0.0031, 32, 16)])
def test_fold_timeseries_bit_exact_vs_oracle(lib, n, period, nbins, nints):
    from oracle import fold as F
    tsamp = 0.00032
    x = pulsar_series(n, tsamp, period, seed=n % 97)
    want = F.fold_series(x, period, tsamp, nbins, nints)
    got = gpu_fold(lib, x, period, tsamp, nbins, nints)
    # both sides add each bin's samples in sample order in float32: identical bits
    assert np.array_equal(got, want)
    assert got.std() > 0


def test_fold_timeseries_rejects_bad_arguments(lib):
    x = torch.zeros(1024, device="cuda")
    assert lib.ps_fold_timeseries(x.data_ptr(), x.data_ptr(), 1024, 16, 0.1, 0.001, 512, None) == -2
    assert lib.ps_fold_timeseries(x.data_ptr(), x.data_ptr(), 8, 16, 0.1, 0.001, 64, None) == -2
    assert lib.ps_fold_timeseries(x.data_ptr(), x.data_ptr(), 1024, 16, 0.0, 0.001, 64, None) == -2
    h = C.c_void_p()
    assert lib.ps_fold_optimiser_create(48, 16, C.byref(h)) == -2 and b"nbins" in lib.ps_last_error()


def gpu_optimise(lib, handle, fold, period, tobs):
    nints, nbins = fold.shape
    d = dev(fold.astype(np.float32).ravel())
    of = np.zeros((nints, nbins), np.float32)
    op = np.zeros(nbins, np.float32)
    res = pb.FoldResult()
    pb.check(lib.ps_fold_optimise(handle, d.data_ptr(), float(period), float(tobs), of.ctypes.data, op.ctypes.data, C.byref(res), None))
    return res, of, op


@pytest.fixture(scope="module")
def optimiser(lib):
    h = C.c_void_p()
    pb.check(lib.ps_fold_optimiser_create(64, 16, C.byref(h)))
    yield h
    lib.ps_fold_optimiser_destroy(h)


@pytest.mark.parametrize("seed,drift,width", [(1, 0.0, 3), (2, 9.0, 6), (3, -14.0, 1), (4, 25.0, 12), (5, -31.0, 20)])
def test_fold_optimise_matches_oracle(lib, optimiser, seed, drift, width):
    """A pulse of `width` bins drifting by `drift` bins across the 16 sub-integrations: the
    optimiser must pick the same template, bin and drift as the oracle and give the same S/N."""
    from oracle import fold as F
    rng = np.random.default_rng(seed)
    fold = rng.normal(0, 1, (16, 64)).astype(np.float32) * 100
    for s in range(16):
        start = int(round(20 + drift * s / 16.0))
        fold[s, (start + np.arange(width)) % 64] += 600
    period, tobs = 0.0731, np.float32(41.94304)
    want = F.FoldOptimiser().optimise(fold, period, tobs)
    res, of, op = gpu_optimise(lib, optimiser, fold, period, tobs)
    assert (res.opt_width, res.opt_bin, res.opt_shift) == (want["opt_width"], want["opt_bin"], want["opt_shift"])
    assert abs(res.opt_shift - (32 - drift)) <= 2          # the drift bank index that undoes the drift
    assert res.peak == pytest.approx(want["peak"], rel=1e-5)
    assert res.opt_sn == pytest.approx(float(want["opt_sn"]), rel=1e-4)
    assert res.sn1 == pytest.approx(float(want["sn"][0]), rel=1e-4) and res.sn2 == pytest.approx(float(want["sn"][1]), rel=1e-4)
    assert np.float32(res.opt_period) == want["opt_period"]
    scale = np.abs(want["opt_fold"]).max()
    assert np.abs(of - want["opt_fold"]).max() <= 1e-5 * scale
    assert np.abs(op - want["opt_prof"]).max() <= 1e-5 * np.abs(want["opt_prof"]).max()


@pytest.mark.parametrize("n,period,nbins,nints", [(1 << 17, 0.25, 32, 16), (1 << 17, 0.0123456, 32, 16), (102400, 0.0817, 16, 16),
                                                  (1 << 20, 1.2345, 32, 8)])
def test_fold_timeseries_equals_reference_kernel(lib, ref, n, period, nbins, nints):
    """Pin against the UNMODIFIED reference kernel run here: fold_time_series_kernel (kernels.cu:597-632,
    forwarded by oracle/ref_harness.cu).  The reference adds with float atomics in no fixed order and
    without a barrier between clearing, adding to and reading its shared bins, so the comparison uses
    inputs on which neither matters: small-integer samples (every partial sum is exact in float32,
    whatever the order) and nbins <= 32, one warp per block, where the missing barrier cannot lose a
    sample; and sub-integrations that are a whole number of nbins-sample strides, because the
    reference's last block otherwise reads up to nbins-1 samples past the end of the series (ours
    clamps there).  On such input the two kernels must agree bit for bit -- bins, hit counts (+1)
    and division."""
    tsamp = 0.00032
    rng = np.random.default_rng(n % 89 + nbins)
    x = rng.integers(0, 16, n).astype(np.float32)
    got = gpu_fold(lib, x, period, tsamp, nbins, nints)
    d_in = dev(x)
    d_out = torch.zeros(nints * nbins, dtype=torch.float32, device="cuda")
    ref.ref_fold_timeseries.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_double, C.c_double, C.c_int]
    assert ref.ref_fold_timeseries(d_in.data_ptr(), d_out.data_ptr(), n, nints, period, tsamp, nbins) == 0, ref.ref_last_error()
    torch.cuda.synchronize()
    want = d_out.cpu().numpy().reshape(nints, nbins)
    assert np.array_equal(got, want)
    assert got.std() > 0


@pytest.mark.parametrize("seed,drift,width", [(1, 0.0, 3), (2, 9.0, 6), (3, -14.0, 1), (4, 25.0, 12), (5, -31.0, 20)])
def test_fold_optimise_equals_reference_optimiser(lib, ref, optimiser, seed, drift, width):
    """Pin against the reference's own FoldOptimiser object (folder.hpp:195-351: its kernels of
    kernels.cu:655-865, three batched cuFFT plans, thrust::max_element) on the same folded
    sub-integrations: same template, bin and optimised period; S/N, profile and optimised
    sub-integrations to transform rounding."""
    rng = np.random.default_rng(seed)
    fold = rng.normal(0, 1, (16, 64)).astype(np.float32) * 100
    for s in range(16):
        start = int(round(20 + drift * s / 16.0))
        fold[s, (start + np.arange(width)) % 64] += 600
    period, tobs = 0.0731, np.float32(41.94304)
    res, of, op = gpu_optimise(lib, optimiser, fold, period, tobs)
    d = dev(fold.ravel())
    sn, per, wid, obin = C.c_float(), C.c_double(), C.c_int(), C.c_int()
    rof = np.zeros((16, 64), np.float32)
    rop = np.zeros(64, np.float32)
    ref.ref_fold_optimise.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_double, C.c_float, C.POINTER(C.c_float),
                                      C.POINTER(C.c_double), C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_void_p, C.c_void_p]
    assert ref.ref_fold_optimise(d.data_ptr(), 64, 16, period, tobs, C.byref(sn), C.byref(per), C.byref(wid), C.byref(obin),
                                 rof.ctypes.data, rop.ctypes.data) == 0, ref.ref_last_error()
    assert (res.opt_width, res.opt_bin) == (wid.value, obin.value)
    assert np.float32(res.opt_period) == np.float32(per.value)      # the reference's getter narrows to float
    assert res.opt_sn == pytest.approx(sn.value, rel=1e-3)
    assert np.abs(of - rof).max() <= 1e-4 * np.abs(rof).max()
    assert np.abs(op - rop).max() <= 1e-4 * np.abs(rop).max()


def test_fold_optimise_is_repeatable(lib, optimiser):
    rng = np.random.default_rng(11)
    fold = rng.normal(0, 1, (16, 64)).astype(np.float32)
    fold[:, 40:44] += 5
    a = gpu_optimise(lib, optimiser, fold, 0.5, 41.9)
    b = gpu_optimise(lib, optimiser, fold, 0.5, 41.9)
    assert a[0].argmax == b[0].argmax and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])


def test_tutorial_fil_with_folding_matches_golden(lib, oracle, tutorial_trials, tmp_path):
    """BASELINE.json configs[0] exactly as the golden run: --npdmp 10 --limit 10.  Final order
    (by max(snr, folded_snr)), folded_snr within 1 %, opt_period to float precision, FOLD blocks
    row by row (see test_oracle_golden for why some golden rows cannot match) -- and the same
    quantities against the oracle's MultiFolder at tight tolerance."""
    from oracle import fold as F
    from test_gpu_filterbank import run_tutorial as run
    recs, dms, _ = run(lib, outdir=str(tmp_path), limit=10, npdmp=10)
    golden = oracle.parse_candidates_peasoup(os.path.join(GOLDEN, "candidates.peasoup"))
    gxml = oracle.parse_overview_candidates(os.path.join(GOLDEN, "overview.xml"))
    written = oracle.parse_candidates_peasoup(os.path.join(tmp_path, "candidates.peasoup"))
    wxml = oracle.parse_overview_candidates(os.path.join(tmp_path, "overview.xml"))
    assert len(written) == len(wxml) == 10
    total = 0
    for i, (w, g, wx, gx) in enumerate(zip(written, golden, wxml, gxml)):
        # same candidate in the same place (acceleration left out: exact S/N ties)
        for k in ("dm_idx", "nh"):
            assert w["dets"][0][k] == g["dets"][0][k], (i, k)
        assert w["dets"][0]["freq"] == pytest.approx(g["dets"][0]["freq"], rel=1e-6)
        assert len(w["dets"]) == len(g["dets"])
        assert wx["snr"] == pytest.approx(gx["snr"], rel=1e-4)
        assert wx["folded_snr"] == pytest.approx(gx["folded_snr"], rel=1e-2), i
        assert wx["opt_period"] == pytest.approx(gx["opt_period"], rel=1e-7), i
        assert w["fold"] is not None and w["fold"].shape == (16, 64)
        n = rows_matching(w["fold"], g["fold"])
        assert n >= 8, (i, n)
        total += n
    assert total >= 110, total
    # against the oracle's restatement of the same stage
    trials, _ = tutorial_trials
    tops = [dict(freq=c["dets"][0]["freq"], dm_idx=c["dets"][0]["dm_idx"], acc=c["dets"][0]["acc"], snr=c["dets"][0]["snr"])
            for c in written]
    res, order = F.fold_candidates(tops, trials, 0.00032, 10)
    assert order == list(range(10))
    for i, (w, wx, r) in enumerate(zip(written, wxml, res)):
        assert wx["folded_snr"] == pytest.approx(float(r["opt_sn"]), rel=1e-3), i
        assert np.float32(wx["opt_period"]) == r["opt_period"], i
        assert np.abs(w["fold"] - r["opt_fold"]).max() <= 1e-4 * np.abs(r["opt_fold"]).max(), i
    # ps_fold_results_last hands the same numbers back through the C ABI
    snr = np.zeros(10, np.float32)
    per = np.zeros(10, np.float64)
    folds = np.zeros((10, 1024), np.float32)
    assert lib.ps_fold_results_last(snr.ctypes.data, per.ctypes.data, folds.ctypes.data, 10) == 10
    assert np.allclose(snr, [x["folded_snr"] for x in wxml], rtol=1e-6)
    assert np.array_equal(folds.reshape(10, 16, 64), np.stack([w["fold"] for w in written]))


def test_npdmp_zero_leaves_candidates_unfolded(lib, oracle, tmp_path):
    run_tutorial(lib, outdir=str(tmp_path), limit=5)
    written = oracle.parse_candidates_peasoup(os.path.join(tmp_path, "candidates.peasoup"))
    assert all(c["fold"] is None for c in written)
    x = oracle.parse_overview_candidates(os.path.join(tmp_path, "overview.xml"))
    assert all(c["folded_snr"] == 0 and c["opt_period"] == 0 for c in x)
