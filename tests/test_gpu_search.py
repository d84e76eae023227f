"""GPU parity of the fused search path (ps_search_*, ps_pipeline_search) through the C ABI:
against the CPU oracle, against the reference's golden candidates.peasoup (BASELINE.json
configs[0]), against the UNMODIFIED reference kernels + cuFFT replaying Worker::start on the same
device (oracle/_ref), and -- at the full 2^23 size -- through domain properties.

Tolerances (north_star): resampling index maps and peak bins bit-exact; spectra within 1e-4
relative; candidate list (DM, accel, bin, nh) identical with S/N within 1 %.  Bins whose value
lies within 1e-3 (relative) of the threshold may legitimately flip because FFT rounding differs
from cuFFT's; they are classified separately and must be rare."""
import ctypes as C
import os

import numpy as np
import pytest

import peasoup_b200 as pb
from peasoup_b200 import ptr
from peasoup_b200.host import make_search_params
from conftest import GOLDEN
from synth import cand_keys, make_trials

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

SPEC_TOL = 1e-4
SNR_TOL = 1e-2


def dev_f32(p, n):
    """copy n floats from a raw device pointer (zero acceleration = identity index map)"""
    t = torch.zeros(n, dtype=torch.float32, device="cuda")
    pb.check(pb.load().ps_resampleII(p, t.data_ptr(), n, 0.0, 1.0, None))
    torch.cuda.synchronize()
    return t.cpu().numpy()


def pipeline_params(size, nsamps, tsamp, nharm=4, min_snr=9.0, max_peaks=0):
    return pb.PipelineParams(make_search_params(size, nsamps, tsamp, nharm, min_snr, max_peaks=max_peaks), 1e-4, 16,
                             0.0, 0.0, 1.1, 64.0, 1400.0, -1.0)


def compare_candidates(got, want, cfg, threshold=9.0):
    """identical (dm_idx, acc, nh, bin, is_top) sets with S/N within 1 %, except records whose S/N is
    within 1e-3 of the threshold on either side (classified, must be < 1 % of the list)."""
    a = cand_keys(got, cfg.nbins, cfg.bin_width_f, with_top=True)
    b = cand_keys(want, cfg.nbins, cfg.bin_width_f, with_top=True)
    only = set(a) ^ set(b)
    marginal = [k for k in only if abs((a.get(k) or b.get(k)) - threshold) < 2e-3 * threshold]
    hard = [k for k in only if k not in marginal]
    assert not hard, (len(a), len(b), hard[:8])
    assert len(marginal) <= max(1, len(b) // 100), marginal
    worst = max((abs(a[k] - b[k]) / b[k] for k in set(a) & set(b)), default=0.0)
    assert worst < SNR_TOL, worst
    return worst, len(marginal)


# ------------------------------------------------------------------ stage-level: whitening + spectra
@pytest.mark.parametrize("size,nsamps,zap", [(1 << 17, (1 << 17) + 300, False), (1 << 17, 100000, True), (1 << 18, 1 << 18, True),
                                             (1 << 20, (1 << 20) + 100, True), (1 << 21, (1 << 21) - 777, False),
                                             (1 << 22, 1 << 22, False)])
def test_whiten_and_spectrum_vs_oracle(lib, oracle, size, nsamps, zap):
    tsamp = 64e-6
    trial = make_trials(1, nsamps, size, tsamp=tsamp, a_inj=20.0, amp=8.0)[0]
    zl = (np.array([50.0, 100.0], np.float32), np.array([0.1, 0.15], np.float32)) if zap else None
    cfg = oracle.SearchConfig(size, tsamp, zap=zl)
    s = pb.Search(make_search_params(size, nsamps, tsamp))
    if zap:
        s.set_zaplist(*zl)
    d_trial = torch.from_numpy(trial).cuda()
    mean, std = s.whiten(d_trial, want_stats=True)
    tim, wmean, wstd, inter = oracle.whiten(trial, cfg, return_intermediates=True)
    got = dev_f32(s.whitened_series_ptr(), size)
    rms = np.sqrt((tim.astype(np.float64) ** 2).mean())
    assert np.abs(got - tim).max() / rms < SPEC_TOL
    assert abs(mean - wmean) / wmean < 1e-5 and abs(std - wstd) / wstd < 1e-5
    for acc in (0.0, 20.0, -35.5):
        s.reset_peaks()
        s.accel_batch([acc])
        p = dev_f32(s.last_spectrum_ptr(), cfg.nbins)
        want = oracle.accel_spectrum(tim, acc, wmean, wstd, cfg)
        # normalised spectrum has unit variance: absolute error == error relative to spectrum rms
        assert np.abs(p - want).max() < 20 * SPEC_TOL  # bins up to ~100 sigma; relative 1e-4 of those
        assert np.sqrt(((p - want) ** 2).mean()) < SPEC_TOL
    s.close()


@pytest.mark.parametrize("size", [1 << 17, 1 << 19, 1 << 20])
def test_raw_peaks_vs_oracle(lib, oracle, size):
    tsamp = 64e-6
    thresh = 6.0  # low threshold: thousands of peaks exercise the compaction
    trial = make_trials(1, size, size, tsamp=tsamp, a_inj=31.4, amp=10.0)[0]
    cfg = oracle.SearchConfig(size, tsamp, min_snr=thresh)
    s = pb.Search(make_search_params(size, size, tsamp, min_snr=thresh))
    s.whiten(torch.from_numpy(trial).cuda())
    accs = np.array([0.0, 31.4, -12.0], np.float32)
    s.reset_peaks()
    s.accel_batch(accs)
    peaks = s.fetch_peaks()
    tim, m, sd = oracle.whiten(trial, cfg)
    want = {}
    for j, acc in enumerate(accs):
        raw = oracle.spectrum_raw_peaks(oracle.accel_spectrum(tim, acc, m, sd, cfg), cfg)
        for nh, (idx, snr, _) in enumerate(raw):
            for i, v in zip(idx, snr):
                want[(j, nh, int(i))] = float(v)
    got = {(int(r["acc_idx"]), int(r["nh"]), int(r["bin"])): float(r["snr"]) for r in peaks}
    assert list(got) == sorted(got)  # (acc_idx, nh, bin) order
    only = set(got) ^ set(want)
    hard = [k for k in only if abs((got.get(k) or want.get(k)) - thresh) > 2e-3 * thresh]
    assert not hard, hard[:5]
    assert len(only) <= max(2, len(want) // 200)
    assert len(want) > 100
    worst = max(abs(got[k] - want[k]) / want[k] for k in set(got) & set(want))
    assert worst < 1e-3
    s.close()


def test_peak_list_overflow_is_reported(lib):
    size, tsamp = 1 << 17, 64e-6
    trial = make_trials(1, size, size, tsamp=tsamp)[0]
    s = pb.Search(make_search_params(size, size, tsamp, min_snr=-100.0, max_peaks=1000))
    s.whiten(torch.from_numpy(trial).cuda())
    s.reset_peaks()
    s.accel_batch([0.0])
    with pytest.raises(pb.PeasoupError, match="max_peaks"):
        s.fetch_peaks()
    s.close()


def test_pipeline_grows_a_full_peak_list(lib, oracle):
    """A DM trial with more above-threshold bins than max_peaks does not fail the pipeline: the list is
    enlarged and the trial's accelerations are searched again; the result equals a run with a
    large list.  (The reference truncates silently at 100000 peaks per spectrum.)"""
    size, tsamp = 1 << 17, 64e-6
    trials = make_trials(2, size, size, tsamp=tsamp, amp=9.0)
    accs = np.array([0.0, 12.0, -30.0], np.float32)
    dms = np.array([3.0, 4.0], np.float32)
    small, _ = pb.pipeline_search(trials, dms, pipeline_params(size, size, tsamp, min_snr=5.0, max_peaks=64), accs=accs)
    large, _ = pb.pipeline_search(trials, dms, pipeline_params(size, size, tsamp, min_snr=5.0, max_peaks=1 << 20), accs=accs)
    assert len(large) > 64 and np.array_equal(small, large)


# ------------------------------------------------------------------ pipeline-level
def test_pipeline_vs_oracle_synthetic(lib, oracle):
    size, nsamps, tsamp = 1 << 17, (1 << 17) + 512, 64e-6
    ndm = 5
    trials = make_trials(ndm, nsamps, size, tsamp=tsamp, a_inj=31.4, amp=9.0)
    dms = np.linspace(0, 40, ndm).astype(np.float32)
    accs = np.array([0.0, -50.0, -25.0, 25.0, 31.4, 50.0], np.float32)
    cfg = oracle.SearchConfig(size, tsamp)
    got, done = pb.pipeline_search(trials, dms, pipeline_params(size, nsamps, tsamp), accs=accs)
    assert done == ndm * len(accs)
    want = oracle.search(trials, dms, accs, cfg)
    assert len(want) > 20
    compare_candidates(got, want, cfg)
    # ascending DM order, a top-level record first
    assert np.all(np.diff(got["dm_idx"]) >= 0) and got["is_top"][0] == 1


@pytest.mark.parametrize("size,nsamps", [(1 << 16, (1 << 16) + 300), (100000, 120000), (1 << 15, 30000)])
def test_pipeline_other_transform_lengths_vs_oracle(lib, oracle, size, nsamps):
    """--fft_size below 2^17 or not a power of two (cuFFT takes any length): the pipeline falls back to
    the reference's statement-by-statement loop on the staged entry points (Worker::start_staged,
    chirp-z transforms) and finds the same candidates as the oracle.  The last case pads with the mean."""
    tsamp = 64e-6
    ndm = 3
    trials = make_trials(ndm, nsamps, size, tsamp=tsamp, a_inj=31.4, amp=12.0)
    dms = np.linspace(0, 20, ndm).astype(np.float32)
    accs = np.array([0.0, -50.0, 31.4], np.float32)
    zl = np.loadtxt(os.path.join(GOLDEN, "default_zaplist.txt"), dtype=np.float32)
    cfg = oracle.SearchConfig(size, tsamp, zap=(zl[:, 0], zl[:, 1]))
    got, done = pb.pipeline_search(trials, dms, pipeline_params(size, nsamps, tsamp), accs=accs, zap=(zl[:, 0], zl[:, 1]))
    assert done == ndm * len(accs)
    want = oracle.search(trials, dms, accs, cfg)
    assert len(want) > 5
    compare_candidates(got, want, cfg)


def test_pipeline_staged_loop_equals_fused_loop(lib, monkeypatch):
    """PS_STAGED=1 sends a power-of-two size through Worker::start_staged: the same records as the
    fused kernels (S/N to 1e-4: the staged loop normalises and sums in separate kernels)."""
    size, nsamps, tsamp = 1 << 17, (1 << 17) + 512, 0.00032
    trials = make_trials(4, nsamps, size, tsamp=tsamp, a_inj=31.4, amp=9.0)
    dms = np.linspace(0, 30, 4).astype(np.float32)
    accs = np.array([0.0, -50.0, 31.4, 50.0], np.float32)
    fused, _ = pb.pipeline_search(trials, dms, pipeline_params(size, nsamps, tsamp), accs=accs)
    monkeypatch.setenv("PS_STAGED", "1")
    staged, done = pb.pipeline_search(trials, dms, pipeline_params(size, nsamps, tsamp), accs=accs)
    assert done == 16 and len(staged) == len(fused) > 20
    for k in ("dm_idx", "acc", "nh", "is_top", "freq"):
        assert np.array_equal(staged[k], fused[k]), k
    assert np.allclose(staged["snr"], fused["snr"], rtol=1e-4)


def test_pipeline_acceleration_plan_and_empty(lib, oracle):
    size, tsamp = 1 << 17, 0.00032
    trials = make_trials(2, size + 10, size, tsamp=tsamp, a_inj=0.0, p0=0.1, amp=6.0, birdies=False)
    prm = pipeline_params(size, size + 10, tsamp)
    prm.acc_start, prm.acc_end, prm.cfreq, prm.foff = -5.0, 5.0, 1475.12, -1.09
    got, done = pb.pipeline_search(trials, [0.0, 5.0], prm)  # AccelerationPlan -> [0,-5,5]
    assert done == 6 and set(got["acc"]) <= {0.0, -5.0, 5.0}
    cfg = oracle.SearchConfig(size, tsamp)
    want = oracle.search(trials, [0.0, 5.0], np.array([0.0, -5.0, 5.0], np.float32), cfg)
    compare_candidates(got, want, cfg)
    got0, done0 = pb.pipeline_search(trials[:0], [], prm)
    assert len(got0) == 0 and done0 == 0


def test_tutorial_fil_matches_golden_candidates(lib, oracle, tutorial_trials):
    """BASELINE.json configs[0] with the parameters the golden run really used."""
    trials, ov = tutorial_trials
    size, tsamp = 1 << 17, 0.00032
    cfg = oracle.SearchConfig(size, tsamp)
    got, done = pb.pipeline_search(trials, ov["dm_list"], pipeline_params(size, trials.shape[1], tsamp), accs=ov["acc_list"])
    assert done == 59 * 3
    gold = np.concatenate([c["dets"] for c in oracle.parse_candidates_peasoup(os.path.join(GOLDEN, "candidates.peasoup"))])
    ours = cand_keys(got, cfg.nbins, cfg.bin_width_f)
    g = cand_keys(gold, cfg.nbins, cfg.bin_width_f)
    missing = [k for k in g if k not in ours]
    assert len(gold) == 1032 and len(g) > 900 and not missing, missing[:5]
    worst = max(abs(ours[k] - g[k]) / g[k] for k in g)
    assert worst < SNR_TOL, worst
    # and the full list equals the oracle's, record for record
    want = oracle.search(trials, ov["dm_list"], ov["acc_list"], cfg)
    compare_candidates(got, want, cfg)
    # DM carried through exactly
    for r in got[:50]:
        assert r["dm"] == ov["dm_list"][r["dm_idx"]]


def ref_search(ref, trials, dms, accs, size, nsamps, tsamp, nharm=4, min_snr=9.0, zapfile=b""):
    class P(C.Structure):
        _fields_ = [("size", C.c_uint), ("nsamps", C.c_uint), ("tsamp", C.c_float), ("nharmonics", C.c_int),
                    ("min_snr", C.c_float), ("min_freq", C.c_float), ("max_freq", C.c_float), ("freq_tol", C.c_float),
                    ("max_harm", C.c_int), ("zapfile", C.c_char_p)]
    p = P(size, nsamps, tsamp, nharm, min_snr, 0.1, 1100.0, 1e-4, 16, zapfile)
    cap = 1 << 20
    o = {k: np.zeros(cap, t) for k, t in [("dm", "f4"), ("dm_idx", "i4"), ("acc", "f4"), ("nh", "i4"), ("snr", "f4"),
                                            ("freq", "f4"), ("is_top", "i4")]}
    dms = np.ascontiguousarray(dms, np.float32)
    accs = np.ascontiguousarray(accs, np.float32)
    secs = C.c_double()
    vp = C.c_void_p
    ref.ref_worker_search.argtypes = [C.POINTER(P), vp, vp, C.c_int, vp, C.c_int, C.c_long] + [vp] * 7 + [C.POINTER(C.c_double)]
    n = ref.ref_worker_search(C.byref(p), ptr(trials), ptr(dms), len(dms), ptr(accs), len(accs), cap,
                              *[ptr(o[k]) for k in ("dm", "dm_idx", "acc", "nh", "snr", "freq", "is_top")], C.byref(secs))
    assert n >= 0, ref.ref_last_error()
    out = np.zeros(n, pb.host.CAND_DTYPE)
    for k in o:
        out[k] = o[k][:n]
    return out, secs.value


def test_pipeline_vs_reference_cuda_build(lib, oracle, ref, tmp_path):
    """Same host trials through the unmodified reference (kernels.cu + transform headers + cuFFT,
    replaying Worker::start) and through ps_pipeline_search: identical candidates."""
    size, nsamps, tsamp = 1 << 18, (1 << 18) - 5000, 64e-6  # padding path
    ndm = 4
    trials = make_trials(ndm, nsamps, size, tsamp=tsamp, a_inj=-20.0, amp=9.0)
    dms = np.linspace(5, 25, ndm).astype(np.float32)
    accs = np.array([0.0, -40.0, -20.0, 20.0], np.float32)
    zf = tmp_path / "zap.txt"
    zf.write_text(open(os.path.join(GOLDEN, "default_zaplist.txt")).read())
    zl = np.loadtxt(zf, dtype=np.float32)
    cfg = oracle.SearchConfig(size, tsamp)
    want, _ = ref_search(ref, trials, dms, accs, size, nsamps, tsamp, zapfile=str(zf).encode())
    got, _ = pb.pipeline_search(trials, dms, pipeline_params(size, nsamps, tsamp), accs=accs, zap=(zl[:, 0], zl[:, 1]))
    assert len(want) > 10
    compare_candidates(got, want, cfg)


@pytest.mark.parametrize("nharm", [0, 1, 2, 3, 5])
def test_pipeline_other_fold_counts_vs_reference_cuda_build(lib, oracle, ref, nharm):
    """-n 0 .. 5 (no harmonic sums ... 32 harmonics) through the whole loop against the unmodified
    reference build: 1-3 folds take the TMA-staged harmonic kernel with fewer ranges, 5 folds the
    gather kernel, 0 folds thresholds the fundamental spectrum alone."""
    size, nsamps, tsamp = 1 << 19, (1 << 19) + 100, 64e-6
    ndm = 2
    trials = make_trials(ndm, nsamps, size, tsamp=tsamp, a_inj=15.0, amp=7.0, seed=nharm)
    dms = np.array([3.0, 9.0], np.float32)
    accs = np.array([0.0, 15.0, -15.0], np.float32)
    cfg = oracle.SearchConfig(size, tsamp, nharmonics=nharm)
    want, _ = ref_search(ref, trials, dms, accs, size, nsamps, tsamp, nharm=nharm)
    got, done = pb.pipeline_search(trials, dms, pipeline_params(size, nsamps, tsamp, nharm=nharm), accs=accs)
    assert done == ndm * len(accs) and len(want) > 0
    compare_candidates(got, want, cfg)
    assert int(got["nh"].max()) <= nharm


def test_tutorial_vs_reference_cuda_build(lib, oracle, ref, tutorial_trials):
    trials, ov = tutorial_trials
    size, tsamp = 1 << 17, 0.00032
    cfg = oracle.SearchConfig(size, tsamp)
    want, _ = ref_search(ref, trials, ov["dm_list"], ov["acc_list"], size, trials.shape[1], tsamp)
    got, _ = pb.pipeline_search(trials, ov["dm_list"], pipeline_params(size, trials.shape[1], tsamp), accs=ov["acc_list"])
    compare_candidates(got, want, cfg)


# ------------------------------------------------------------------ full size (BASELINE configs[1])
def test_full_size_injected_pulsar_properties(lib, oracle):
    size, tsamp = 1 << 23, 64e-6
    nsamps = size + 4096
    a_inj, p0 = 31.4, 7.5e-3
    trials = make_trials(1, nsamps, size, tsamp=tsamp, a_inj=a_inj, p0=p0, amp=3.0, seed=7)
    accs = np.arange(-50, 51, 10, dtype=np.float32)
    accs = np.sort(np.append(accs, np.float32(a_inj)))
    zl = np.loadtxt(os.path.join(GOLDEN, "default_zaplist.txt"), dtype=np.float32)
    got, done = pb.pipeline_search(trials, [42.0], pipeline_params(size, nsamps, tsamp), accs=accs, zap=(zl[:, 0], zl[:, 1]))
    assert done == len(accs)
    top = got[got["is_top"] == 1]
    best = top[np.argmax(top["snr"])]
    # the strongest candidate is the injected pulsar: at the injected acceleration, at a harmonic
    # of 1/p0 (a 5 % duty pulse train peaks in a summed spectrum), well above threshold
    assert best["acc"] == np.float32(a_inj), best
    ratio = best["freq"] * p0
    assert abs(ratio - round(ratio)) < 2e-3 or abs(1 / ratio - round(1 / ratio)) < 2e-3, best
    assert best["snr"] > 30
    # zapped birdies (50/100 Hz) produce no candidates
    assert not np.any((np.abs(got["freq"] - 50.0) < 0.05) | (np.abs(got["freq"] - 100.0) < 0.05))
    # the acceleration distiller keeps the neighbouring-acceleration detections as associates
    assert (got["is_top"] == 0).sum() > 0


@pytest.mark.parametrize("size,nbatch,pad,nharm", [(1 << 17, 3, 7, 4), (1 << 20, 2, 7, 4), (1 << 17, 3, 64, 4), (1 << 20, 2, 8, 4),
                                                   (1 << 18, 2, 12, 3), (1 << 18, 2, 4, 2), (1 << 18, 2, 4, 1), (1 << 17, 2, 7, 5)])
def test_harmonic_peaks_stage_bit_exact_vs_oracle(lib, oracle, size, nbatch, pad, nharm):
    """ps_search_harmonic_peaks on spectra supplied by the caller (BASELINE.json configs[2]'s stage):
    same bins, same levels and bit-identical S/N as harmonic_sum_kernel + device_find_peaks."""
    tsamp, thresh = 64e-6, 5.0
    cfg = oracle.SearchConfig(size, tsamp, nharmonics=nharm, min_snr=thresh)
    rng = np.random.default_rng(size + nbatch)
    # spectra need not be packed: an odd stride takes the gather kernel, a stride that keeps every
    # spectrum 16-byte aligned with >= 4 floats of padding takes the TMA-staged kernel
    stride = cfg.nbins + pad
    stride += (-stride) % 4 if pad % 4 == 0 else 0
    spectra = np.zeros((nbatch, stride), np.float32)
    for b in range(nbatch):
        p = rng.normal(0, 1, cfg.nbins).astype(np.float32)
        p[::32] += 1.0                                        # harmonic_sum_test.cpp's pattern
        f0 = 1000 + 37 * b
        p[f0 * np.arange(1, 17)] += 6.0                       # comb: fundamental + 15 harmonics
        p[[2047, 2048, 4095, 4096, cfg.nbins - 2, cfg.nbins - 1]] += 40.0   # tile edges of the staged kernel, last bins
        spectra[b, :cfg.nbins] = p
    s = pb.Search(make_search_params(size, size, tsamp, nharm, min_snr=thresh))
    s.reset_peaks()
    d = torch.from_numpy(spectra).cuda()
    s.harmonic_peaks(d, stride, nbatch, acc_idx0=10)
    peaks = s.fetch_peaks()
    want = {}
    for b in range(nbatch):
        for nh, (idx, snr, _) in enumerate(oracle.spectrum_raw_peaks(spectra[b, :cfg.nbins], cfg)):
            for i, v in zip(idx, snr):
                want[(10 + b, nh, int(i))] = np.float32(v)
    got = {(int(r["acc_idx"]), int(r["nh"]), int(r["bin"])): np.float32(r["snr"]) for r in peaks}
    assert len(want) > 50 and set(got) == set(want)
    assert all(got[k] == want[k] for k in want)
    if nharm >= 4:
        assert any(k[1] == 4 and k[2] == 1000 * 16 for k in got)  # the comb is found at the top level
    assert any(k[2] == 2048 for k in got)
    if nharm >= 4:
        assert any(k[2] == cfg.nbins - 1 for k in got)   # the last bin, alone in the staged kernel's last tile
    s.close()


@pytest.mark.parametrize("size,acc_hi,a_inj", [(1 << 23, 50.0, 31.4), (1 << 24, 500.0, 312.5), (1 << 25, 500.0, 312.5)])
def test_survey_sizes_vs_reference_cuda_build(lib, oracle, ref, size, acc_hi, a_inj, tmp_path):
    """BASELINE.json configs[3] and [4] shapes (2^23 samples, +-50 m/s^2; 2^24 samples, +-500 m/s^2)
    on one DM trial with the FULL acceleration lists of those configurations (101 and 201 trials):
    identical candidates to the unmodified reference build.  At 2^24 samples and 500 m/s^2 the resampling shift reaches ~3700 samples and
    crosses a rounding boundary far more often than in the other tests (FP64 fallback of the loader).
    2^25 samples is the largest transform the engine takes; the loader's float32 shift estimate is
    taken at even sample indices only, which are exact in float32 up to 2^25 (the index map of the
    product loader is pinned bit for bit in test_product_resample_index_map_bit_exact)."""
    tsamp = 64e-6
    nsamps = size + 4096
    trials = make_trials(1, nsamps, size, tsamp=tsamp, a_inj=a_inj, amp=3.0, seed=size % 1000)
    step = 1.0 if acc_hi == 50.0 else 5.0
    accs = np.sort(np.append(np.arange(-acc_hi, acc_hi + step / 2, step), a_inj)).astype(np.float32)
    assert len(accs) in (102, 202)
    dms = np.array([12.5], np.float32)
    cfg = oracle.SearchConfig(size, tsamp)
    zf = tmp_path / "zap.txt"
    zf.write_text(open(os.path.join(GOLDEN, "default_zaplist.txt")).read())
    zl = np.loadtxt(zf, dtype=np.float32)
    want, _ = ref_search(ref, trials, dms, accs, size, nsamps, tsamp, zapfile=str(zf).encode())
    got, done = pb.pipeline_search(trials, dms, pipeline_params(size, nsamps, tsamp, max_peaks=1 << 22), accs=accs,
                                   zap=(zl[:, 0], zl[:, 1]))
    assert done == len(accs) and len(want) > 5
    compare_candidates(got, want, cfg)
    assert np.any((got["acc"] == np.float32(a_inj)) & (got["snr"] > 20))  # the injected pulsar, at its acceleration


def test_two_workers_on_two_gpus_equal_one(lib):
    """`-t 2` of the reference (one host thread per GPU, mutex-protected DMDispenser,
    pipeline_multi.cu:342-359): the merged candidate list equals the one-worker list."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    size, tsamp = 1 << 18, 64e-6
    ndm = 12
    trials = make_trials(ndm, size, size, tsamp=tsamp, a_inj=-20.0, amp=9.0)
    dms = np.linspace(0, 60, ndm).astype(np.float32)
    accs = np.array([0.0, -20.0, 15.0], np.float32)
    one, d1 = pb.pipeline_search(trials, dms, pipeline_params(size, size, tsamp), accs=accs, nworkers=1)
    two, d2 = pb.pipeline_search(trials, dms, pipeline_params(size, size, tsamp), accs=accs, nworkers=2)
    assert d1 == d2 == ndm * len(accs) and len(one) > 10
    assert np.array_equal(one, two)   # ps_pipeline_search returns ascending dm_idx order


def test_fused_whitening_zaps_the_47tuc_list(lib, oracle):
    """The fused whitening marks zapped bins on the median, one warp per birdie (zap_mark_kernel), and
    writes (1, 0) there in the de-reddening pass.  With the reference's 104-line survey zaplist
    (tests/golden/47tuc.zaplist): transform the whitened series back and check that EXACTLY the bins
    the reference's zap_birdies_kernel arithmetic selects (oracle.zap, pinned against the reference
    kernel in test_zap_47tuc_zaplist_vs_reference) hold (1, 0); and the series itself against the
    oracle's whitening with the same list."""
    size, tsamp = 1 << 20, 64e-6
    zl = np.loadtxt(os.path.join(GOLDEN, "47tuc.zaplist"), dtype=np.float32)
    trial = make_trials(1, size, size, tsamp=tsamp, a_inj=20.0, amp=8.0)[0]
    cfg = oracle.SearchConfig(size, tsamp, zap=(zl[:, 0], zl[:, 1]))
    s = pb.Search(make_search_params(size, size, tsamp))
    s.set_zaplist(zl[:, 0], zl[:, 1])
    s.whiten(torch.from_numpy(trial).cuda())
    got = dev_f32(s.whitened_series_ptr(), size).copy()
    tim, _, _ = oracle.whiten(trial, cfg)
    assert np.abs(got - tim).max() / np.sqrt((tim.astype(np.float64) ** 2).mean()) < SPEC_TOL
    marked = oracle.zap(np.zeros(cfg.nbins, np.complex64), zl[:, 0], zl[:, 1], cfg.bin_width_f) == 1
    assert marked.sum() > 104
    Y = np.fft.rfft(got.astype(np.float64)) / size          # C2R is unnormalised: back to the whitened bins
    is_one = np.abs(Y - 1.0) < 1e-3
    assert is_one[marked].all()
    assert (is_one & ~marked).sum() <= 3                   # a whitened noise bin lands that close to (1, 0) once in ~10^6
    s.close()


# ------------------------------------------------------------------ the product gather's index map
def _map_cases():
    c = []
    for a in np.arange(-50.0, 50.5, 1.0):
        c.append((1 << 23, float(a)))
    for a in np.arange(-500.0, 502.5, 5.0):
        c.append((1 << 24, float(a)))
    for a in (-500.0, 500.0, 312.5, -0.5):
        c.append((1 << 25, a))
    for a in (-50.0, 50.0, 7.0):
        c += [(1 << 17, a * 10), (1 << 19, a), (1 << 20, a), (1 << 21, a), (1 << 22, a)]
    return c


@pytest.mark.parametrize("size", [1 << 17, 1 << 19, 1 << 20, 1 << 21, 1 << 22, 1 << 23, 1 << 24, 1 << 25])
def test_product_resample_index_map_bit_exact(lib, oracle, size):
    """north_star: resampling index maps are bit-exact.  ps_search_resample_index_map runs the loader
    that ps_search_accel_batch itself uses (fft.cuh LoadResampled, set up the same way) over a series
    whose sample i holds the integer i, so what comes back is the index each output sample was read
    from.  Every acceleration of the survey lists (2^23: -50..50 step 1; 2^24: -500..500 step 5;
    2^25: the extremes) against resample_kernelII's FP64 expression (kernels.cu:314-317) in the
    oracle; and the share of sample pairs that needed the FP64 expression stays small."""
    tsamp = 64e-6
    s = pb.Search(make_search_params(size, size, tsamp))
    d_idx = torch.zeros(size, dtype=torch.int32, device="cuda")
    accs = [a for (n, a) in _map_cases() if n == size]
    assert accs
    worst = 0.0
    for a in accs:
        want = oracle.resample_index(size, a, tsamp, 2)
        d_idx.zero_()
        exact = s.resample_index_map(a, d_idx)
        got = d_idx.cpu().numpy().view(np.uint32).astype(np.uint64)
        assert np.array_equal(got, want), (size, a, int(np.flatnonzero(got != want)[0]))
        worst = max(worst, exact / (size / 2))
    assert worst < 0.03, (size, worst)   # float32 estimate decides > 97 % of the pairs even at 2^25 / 500 m/s^2
    s.close()


_BAND_SCRIPT = r"""
import sys, numpy as np, torch
sys.path.insert(0, {repo!r}); sys.path.insert(0, {tests!r})
import peasoup_b200 as pb
from peasoup_b200.host import make_search_params
from synth import make_trials
size, tsamp = 1 << 19, 64e-6
trial = make_trials(1, size, size, tsamp=tsamp, a_inj=31.4, amp=10.0)[0]
s = pb.Search(make_search_params(size, size, tsamp, min_snr=6.0))
s.whiten(torch.from_numpy(trial).cuda())
accs = np.array([0.0, 31.4, -12.0, 400.0, 401.0, -350.0, 5.0], np.float32)
s.reset_peaks(); s.accel_batch(accs)
np.save(sys.argv[1], s.fetch_peaks())
"""


def test_pass1_acceleration_groups_equal_one_group(lib, tmp_path):
    """ps_search_accel_batch runs pass 1 of a launch in groups of neighbouring accelerations (PS_BAND_MB of
    shift band per group, each group into its own slices of the inter-pass buffer).  With a band of a few
    bytes every trial is a group of its own, with a huge one the launch is one group: same raw peaks."""
    import subprocess
    import sys
    script = tmp_path / "band.py"
    script.write_text(_BAND_SCRIPT.format(repo=pb.REPO, tests=os.path.join(pb.REPO, "tests")))
    out = {}
    for name, mb in (("split", "0.000001"), ("whole", "1000000")):
        env = dict(os.environ, PS_BAND_MB=mb)
        path = tmp_path / f"{name}.npy"
        subprocess.run([sys.executable, str(script), str(path)], check=True, env=env, timeout=300)
        out[name] = np.load(path)
    assert len(out["split"]) > 100
    assert out["split"].tobytes() == out["whole"].tobytes()
