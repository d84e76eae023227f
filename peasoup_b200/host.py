"""Thin Python conveniences over the C ABI, used by tests/ and bench.py.

Nothing here computes: each function marshals numpy arrays / torch tensors into one call of
libpeasoup_b200.so.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import (Candidate, PipelineParams, RawPeak, SearchParams, check, load, ptr)

CAND_DTYPE = np.dtype([("dm", "f4"), ("dm_idx", "i4"), ("acc", "f4"), ("nh", "i4"), ("snr", "f4"),
                       ("freq", "f4"), ("is_top", "i4")])
PEAK_DTYPE = np.dtype([("acc_idx", "i4"), ("nh", "i4"), ("bin", "i4"), ("snr", "f4")])
assert CAND_DTYPE.itemsize == C.sizeof(Candidate) and PEAK_DTYPE.itemsize == C.sizeof(RawPeak)


def stream_ptr(stream=None) -> int:
    if stream is None:
        return 0
    return int(stream.cuda_stream)


class FFTPlan:
    """ps_fft_plan (replaces CuFFTerR2C / CuFFTerC2R, ffter.hpp:46-78)."""

    def __init__(self, n: int):
        self.n = n
        self._h = C.c_void_p()
        check(load().ps_fft_plan_create(n, C.byref(self._h)))

    def r2c(self, d_in, d_out, stream=None):
        check(load().ps_fft_r2c(self._h, ptr(d_in), ptr(d_out), stream_ptr(stream)))

    def c2r(self, d_in, d_out, stream=None):
        check(load().ps_fft_c2r(self._h, ptr(d_in), ptr(d_out), stream_ptr(stream)))

    def close(self):
        if self._h:
            load().ps_fft_plan_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def make_search_params(size, nsamps, tsamp, nharmonics=4, min_snr=9.0, min_freq=0.1, max_freq=1100.0,
                       max_peaks=0) -> SearchParams:
    return SearchParams(size, nsamps, tsamp, nharmonics, min_snr, min_freq, max_freq, 0.05, 0.5, max_peaks)


class Search:
    """ps_search handle: the fused per-DM device path (Worker::start, pipeline_multi.cu:100-252)."""

    def __init__(self, params: SearchParams):
        self.params = params
        self._h = C.c_void_p()
        check(load().ps_search_create(C.byref(params), C.byref(self._h)))
        self._peaks = np.zeros(params.max_peaks or (1 << 20), PEAK_DTYPE)

    def set_zaplist(self, freqs, widths):
        f = np.ascontiguousarray(freqs, np.float32)
        w = np.ascontiguousarray(widths, np.float32)
        check(load().ps_search_set_zaplist(self._h, ptr(f), ptr(w), len(f)))

    def whiten(self, d_trial_u8, stream=None, want_stats=False):
        if want_stats:
            m, s = C.c_float(), C.c_float()
            check(load().ps_search_whiten(self._h, ptr(d_trial_u8), C.byref(m), C.byref(s), stream_ptr(stream)))
            return m.value, s.value
        check(load().ps_search_whiten(self._h, ptr(d_trial_u8), None, None, stream_ptr(stream)))

    def accel_batch(self, accs, acc_idx0=0, stream=None):
        a = np.ascontiguousarray(accs, np.float32)
        check(load().ps_search_accel_batch(self._h, ptr(a), len(a), acc_idx0, stream_ptr(stream)))

    def harmonic_peaks(self, d_spectra, stride, nbatch, acc_idx0=0, stream=None):
        """Harmonic sums + peaks of `nbatch` normalised spectra already on the device."""
        check(load().ps_search_harmonic_peaks(self._h, ptr(d_spectra), stride, nbatch, acc_idx0, stream_ptr(stream)))

    def reset_peaks(self, stream=None):
        check(load().ps_search_reset_peaks(self._h, stream_ptr(stream)))

    def fetch_peaks(self, stream=None) -> np.ndarray:
        n = check(load().ps_search_fetch_peaks(self._h, ptr(self._peaks), len(self._peaks), stream_ptr(stream)))
        return self._peaks[:n].copy()

    def resample_index_map(self, acc, d_idx, stream=None) -> int:
        """Index map of the search path's resampling gather into d_idx (uint32, size); returns the
        number of sample pairs that took the FP64 path."""
        n = C.c_uint()
        check(load().ps_search_resample_index_map(self._h, float(acc), ptr(d_idx), C.byref(n), stream_ptr(stream)))
        return n.value

    def whitened_series_ptr(self) -> int:
        return load().ps_search_whitened_series(self._h)

    def last_spectrum_ptr(self) -> int:
        return load().ps_search_last_spectrum(self._h)

    def close(self):
        if self._h:
            load().ps_search_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def accel_plan(acc_lo, acc_hi, tol, pulse_width, nsamps, tsamp, cfreq, bw, dm) -> np.ndarray:
    """AccelerationPlan::generate_accel_list (utils.hpp:168-192)."""
    out = np.zeros(1 << 16, np.float32)
    n = load().ps_accel_plan_generate(acc_lo, acc_hi, tol, pulse_width, nsamps, tsamp, cfreq, bw, dm, ptr(out), len(out))
    if n > len(out):
        out = np.zeros(n, np.float32)
        load().ps_accel_plan_generate(acc_lo, acc_hi, tol, pulse_width, nsamps, tsamp, cfreq, bw, dm, ptr(out), n)
    return out[:n].copy()


def cluster_peaks(idxs, snrs, min_gap=30):
    i = np.ascontiguousarray(idxs, np.int32)
    s = np.ascontiguousarray(snrs, np.float32)
    oi, os_ = np.zeros_like(i), np.zeros_like(s)
    n = load().ps_cluster_peaks(ptr(i), ptr(s), len(i), min_gap, ptr(oi), ptr(os_))
    return oi[:n], os_[:n]


def peak_range(nbins, bin_width, nh, min_freq, max_freq):
    a, b, f = C.c_int(), C.c_int(), C.c_float()
    check(load().ps_peak_range(nbins, bin_width, nh, min_freq, max_freq, C.byref(a), C.byref(b), C.byref(f)))
    return a.value, b.value, f.value


def distill(kind, cands: np.ndarray, tol, max_harm_or_tobs, keep_related, fractional=True) -> np.ndarray:
    kinds = {"harmonic": 0, "acceleration": 1, "dm": 2}
    c = np.ascontiguousarray(cands, CAND_DTYPE)
    cap = max(16, 4 * len(c) + 16)
    while True:
        out = np.zeros(cap, CAND_DTYPE)
        n = check(load().ps_distill(kinds[kind], tol, max_harm_or_tobs, int(keep_related), int(fractional), ptr(c),
                                    len(c), ptr(out), cap))
        if n <= cap:
            return out[:n].copy()
        cap = n


def distill_merged(recs: np.ndarray, freq_tol=1e-4, max_harm=16) -> np.ndarray:
    """DMDistiller + HarmonicDistiller on the merged per-DM list (pipeline_multi.cu:353-365)."""
    c = np.ascontiguousarray(recs, CAND_DTYPE)
    cap = max(16, len(c) + 16)
    while True:
        out = np.zeros(cap, CAND_DTYPE)
        n = check(load().ps_distill_merged(ptr(c), len(c), freq_tol, int(max_harm), ptr(out), cap))
        if n <= cap:
            return out[:n].copy()
        cap = n


def pipeline_search(trials_u8, dm_list, params: PipelineParams, accs=None, zap=None, device0=0, nworkers=1,
                    dm_idx_offset=0, max_out=1 << 20):
    """ps_pipeline_search: host u8 trials [ndm][nsamps] -> flattened candidate records.

    trials_u8: numpy uint8 array or (pinned) torch CPU tensor.  Returns (records, trials_done)."""
    dm = np.ascontiguousarray(dm_list, np.float32)
    a = np.ascontiguousarray(accs if accs is not None else [], np.float32)
    zf = np.ascontiguousarray(zap[0] if zap else [], np.float32)
    zw = np.ascontiguousarray(zap[1] if zap else [], np.float32)
    out = np.zeros(max_out, CAND_DTYPE)
    done = C.c_ulonglong(0)
    n = check(load().ps_pipeline_search(C.byref(params), ptr(trials_u8), ptr(dm), len(dm), dm_idx_offset, ptr(a), len(a),
                                        ptr(zf), ptr(zw), len(zf), device0, nworkers, ptr(out), max_out, C.byref(done)))
    if n > max_out:
        raise RuntimeError(f"candidate buffer too small: {n} > {max_out}")
    return out[:n].copy(), done.value


class Pipeline:
    """ps_pipeline handle: worker device state built once, run() per batch of DM trials."""

    def __init__(self, params: PipelineParams, zap=None, device0=0, nworkers=1, max_out=1 << 20):
        zf = np.ascontiguousarray(zap[0] if zap else [], np.float32)
        zw = np.ascontiguousarray(zap[1] if zap else [], np.float32)
        self._h = C.c_void_p()
        self._out = np.zeros(max_out, CAND_DTYPE)
        check(load().ps_pipeline_open(C.byref(params), ptr(zf), ptr(zw), len(zf), device0, nworkers, C.byref(self._h)))

    def run(self, trials_u8, dm_list, accs=None, dm_idx_offset=0):
        dm = np.ascontiguousarray(dm_list, np.float32)
        a = np.ascontiguousarray(accs if accs is not None else [], np.float32)
        done = C.c_ulonglong(0)
        n = check(load().ps_pipeline_run(self._h, ptr(trials_u8), ptr(dm), len(dm), dm_idx_offset, ptr(a), len(a),
                                         ptr(self._out), len(self._out), C.byref(done)))
        if n > len(self._out):
            raise RuntimeError(f"candidate buffer too small: {n} > {len(self._out)}")
        return self._out[:n].copy(), done.value

    def close(self):
        if self._h:
            load().ps_pipeline_close(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
